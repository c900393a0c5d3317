export MT_IGEMM_DBG=32
for W in 0 1; do echo "=== MT_HALO_WIDE=$W"; MT_HALO_WIDE=$W MT_SHAPE=256,256,3,1,66,conv MT_STATS=1 timeout 300 python tools/conv_timeline.py 2>&1 | grep "back-to-back\|tile [0-3]:\|setup\|kernel "; done
