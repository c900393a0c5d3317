# timelines of CTA 0 for the narrow conv classes (what bounds them: MMA, operand stream or epilogue)
export MT_IGEMM_DBG=32
run() { echo "=== $1 stats=$2"; MT_SHAPE=$1 MT_STATS=$2 timeout 300 python tools/conv_timeline.py 2>&1 | tail -8; }
run 128,64,3,2,128,convT ""
run 128,64,3,2,128,convT 1
run 64,128,3,2,258,dgrad ""
run 3,64,7,1,262,stem 1
run 64,128,3,1,130,dgrad ""
run 256,256,3,1,66,conv 1
