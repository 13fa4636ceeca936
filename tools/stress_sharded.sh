# hold a context on GPU 0 in another process (as pytest's parent does), then run the check repeatedly
python -c "
import linearprogramming_b200 as lp, time
c=lp.Context(0); time.sleep(100)" &
HOLD=$!
sleep 8
for cfg in "p2p 9" "p2p 0" "nccl 0" "p2p 9" "p2p 0"; do set -- $cfg
  MP_TRANSPORTS=$1 MP_WINDOW=$2 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29557 tests/mp_sharded_check.py > gpurun_out/stress_$1_$2.log 2>&1
  echo "$cfg rc=$?" >> gpurun_out/stress.log
  grep -h "AssertionError: (" gpurun_out/stress_$1_$2.log | cut -c1-300 >> gpurun_out/stress.log
done
kill $HOLD
