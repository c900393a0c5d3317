"""The number to beat on the same GPU (SURVEY 8d "GPU baseline in the same run"): the reference's
algorithm as eager PyTorch fp32 on the B200 — oracle/step.py (pinned against the unmodified
reference, tests/test_oracle_golden.py) with its tensors moved to CUDA; noise / eps / z drawn on the
CPU and copied, as the reference does (misc.py:25, networks.py:137, adain_model.py:84).
/root/reference itself does not exist on the GPU box.  Prints images/s for true fp32 and TF32."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench
from oracle import step as ostep, synth

B = int(os.environ.get("MT_B", "8"))
meta = json.loads(str(np.load(os.path.join(ROOT, "tests", "golden", "nets_dim64_256.npz"))["meta"]))
sh = meta["shapes"]
sh = {"content_encoder": sh["content_encoder"], "style_encoder": sh["style_encoder"],
      "decoder": sh["decoder"], "discriminator1": sh["discriminator"], "discriminator2": sh["discriminator"]}


class CudaDraws:
    def __init__(self, seed):
        self.d = synth.Draws(seed)

    def __call__(self, shape):
        return self.d(shape).cuda(non_blocking=True)


for tf32 in (False, True):
    torch.backends.cudnn.allow_tf32 = tf32
    torch.backends.cuda.matmul.allow_tf32 = tf32
    torch.backends.cudnn.benchmark = True
    params = {n: {k: v.cuda() for k, v in synth.synth_params(s_, 0, n + ".").items()} for n, s_ in sh.items()}
    st = ostep.AdaINStep(params, B)
    b = bench.synth_batch(B, 0, False)
    img, c = torch.cat((b["x1"], b["x2"])).cuda(), torch.cat((b["y1"], b["y2"])).cuda()
    draws = CudaDraws(0)
    for _ in range(3):
        st.optimize_parameters(img, c, draws)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    K = 5
    for _ in range(K):
        st.optimize_parameters(img, c, draws)
    torch.cuda.synchronize()
    sec = (time.perf_counter() - t0) / K
    print(json.dumps({"impl": "eager torch port of the reference on cuda:0", "tf32": tf32, "batch": B,
                      "ms_per_iter": sec * 1e3, "images_per_sec": 2 * B / sec}), flush=True)
    del st, params
    torch.cuda.empty_cache()
