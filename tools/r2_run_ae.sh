for S in 256,256,3,1,66 64,128,3,2,258 128,256,3,2,130; do
for D in 0 1 2 3 4 7; do MT_SHAPE=$S MT_IGEMM_DBG=$D timeout 300 python tools/wgrad_time.py 2>/dev/null | tail -1; done; done
