"""Training-step throughput (images/s): the drop-in UNet in train() mode (our forward + backward kernels, torch's loss
and Adam) against the reference module on the same GPU through PyTorch/cuDNN autograd (fp32 as the reference trains,
and bf16 autocast).  `python tools/train_bench.py > profiles/r02_train_bench.json`   (measurement aid)"""
import importlib.util, json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "robotic-discovery-platform_b200")]
from oracle.make_golden_train import train_batch
from pkg.segmentation_model import UNet
from rdp_b200 import synth

def ref_module():
    path = os.path.join(ROOT, "oracle", "_ref", "pkg", "segmentation_model.py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("_rdp_ref_seg_train", path)
    m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
    return m

def timed_steps(model, x, t, autocast=None, steps=8):
    opt = torch.optim.Adam(model.parameters(), lr=1e-4)
    crit = torch.nn.BCEWithLogitsLoss()
    def step():
        opt.zero_grad()
        if autocast is not None:
            with torch.autocast("cuda", dtype=autocast):
                out = model(x)
            loss = crit(out.float(), t)
        else:
            loss = crit(model(x), t)
        loss.backward()
        opt.step()
        return loss
    for _ in range(3): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps): loss = step()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, float(loss)

out = {}
seg = ref_module()
for B in (4, 32):
    x, t = (torch.from_numpy(a).cuda() for a in train_batch(tuple(range(40, 40 + B)), 256, 256))
    row = {}
    ours = UNet(3, 1); ours.load_state_dict(synth.make_state_dict(7, planted=False)); ours = ours.cuda().train()
    ms, loss = timed_steps(ours, x, t)
    row["ours"] = {"ms_per_step": ms, "images_per_s": B / ms * 1e3, "loss_after": loss}
    # forward + backward GFLOP: 3 x forward conv FLOPs (forward, input gradient, weight gradient), 79.96 GFLOP per 256x256 frame
    row["ours"]["tflops_equiv"] = 3 * 79.960e9 * B / (ms * 1e-3) / 1e12
    del ours
    if seg is not None:
        for name, ac in (("reference_cudnn_fp32", None), ("reference_cudnn_bf16_autocast", torch.bfloat16)):
            torch.backends.cudnn.benchmark = True
            ref = seg.UNet(3, 1); ref.load_state_dict(synth.make_state_dict(7, planted=False)); ref = ref.cuda().train()
            ms, loss = timed_steps(ref, x, t, ac)
            row[name] = {"ms_per_step": ms, "images_per_s": B / ms * 1e3, "loss_after": loss}
            del ref
    torch.cuda.empty_cache()
    out[f"batch{B}_256x256"] = row
print(json.dumps(out, indent=1))
