timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q 2>&1 | tail -3
export MT_IGEMM_DBG=32
run() { echo "=== $1 stats=$2"; MT_SHAPE=$1 MT_STATS=$2 timeout 300 python tools/conv_timeline.py 2>&1 | grep "back-to-back\|tile [2-3]:\|epilogue of"; }
run 64,128,3,1,130,conv 1
run 256,256,3,1,66,conv 1
unset MT_IGEMM_DBG
for i in 1 2 3; do timeout 600 python bench.py --steps 20 --warmup 5 --no-baselines 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks']['sm_mhz'], d['roofline']['achieved'])"; done
