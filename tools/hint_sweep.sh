# L2 priority hints and ring task size beside the steps (8192 x 16384, default path)
for h in 0 1 2 3; do echo "== LPGPU_L2HINT=$h"; LPGPU_L2HINT=$h BLK_VARIANTS="default=0x0" timeout 60 python tools/blocked_time.py 8192 1312 2>&1 | grep -v "^rank1\|log equal"; done
for rb in 4 16; do echo "== LPGPU_FL_RB=$rb"; LPGPU_FL_RB=$rb BLK_VARIANTS="default=0x0" timeout 60 python tools/blocked_time.py 8192 1312 2>&1 | grep -v "^rank1\|log equal"; done
