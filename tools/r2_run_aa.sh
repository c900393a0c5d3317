for B in 0 128 64; do echo "MT_IGEMM_BN=$B"; MT_IGEMM_BN=$B timeout 600 python tools/ncu_conv_classes.py 2>/dev/null | grep "^D 3x3" | grep -v wgrad; done
