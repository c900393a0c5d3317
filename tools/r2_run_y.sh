export MT_IGEMM_DBG=32
run() { echo "=== $1 stats=$2 halo=$3"; MT_HALO=$3 MT_SHAPE=$1 MT_STATS=$2 timeout 300 python tools/conv_timeline.py 2>&1 | grep "back-to-back\|tile [2-5]:\|setup"; }
run 64,128,3,1,130,conv 1 1
run 64,128,3,1,130,conv 1 0
run 64,64,3,1,130,conv 1 1
run 128,64,3,2,128,convT 1 1
run 128,128,3,1,66,conv 1 1
