# ring depth sweep: pipelined (ring pass) and in place, 8192 x 16384
for nd in 2 3 4 5; do echo "== LPGPU_RING_ND=$nd"; LPGPU_RING_ND=$nd BLK_VARIANTS="ring_pipe32=0x201000,ring_pipe28=0x1c1000,ring_inpl=0x2000" timeout 120 python tools/blocked_time.py 8192 1312 2>&1 | grep -v "^rank1\|log equal"; done
