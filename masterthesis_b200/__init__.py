"""masterthesis_b200 — B200-native hot path of kartikkadur/MasterThesis (translation GAN nets).

Layout:
  csrc/      hand-written sm_100a CUDA (tcgen05/TMEM/TMA implicit-GEMM convs, norm/elementwise)
             behind the C ABI of include/mt_b200.h -> libmt_b200.so
  ops.py     tensor-level wrappers over that ABI        functional.py  autograd block plans
  models/    host-side mirror of the reference's `models` package (same class names / API)
  optim.py   flat-bucket fused Adam                      parallel.py    one-process-per-GPU DP
"""
__version__ = "0.1.0"
