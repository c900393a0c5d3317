for D in 0 4 7; do echo "MT_IGEMM_DBG=$D"; MT_IGEMM_DBG=$D timeout 600 python tools/ncu_conv_classes.py 2>/dev/null | grep "wgrad" | grep "^res 3x3\|^down 3x3 s2 64\|^D 3x3" ; done
