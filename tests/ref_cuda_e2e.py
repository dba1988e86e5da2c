"""GPU-side end-to-end comparator (test-side tool; not the bench's reference arm, not product code).

What the reference would do on this B200: its own module structure (this repo's host glue, which mirrors
models/aff_transformer.py line by line) driven by an operator set made of
  * the REFERENCE's compiled CLUSTEN kernels (oracle/_ref, built unmodified by oracle/Makefile) behind autograd
    Functions with the call order of clusten/clusten.py:17-92,
  * stock torch ops for everything else, exactly the expressions of the reference (oracle/aff_ops_ref.py,
    oracle/point_ref.py evaluated on the GPU),
  * kNN: pykeops is not installed here; this repo's affb200_knn2d stands in for it (that can only favour the comparator),
eager launches, fp16 autocast + GradScaler as the reference trains (config.py:178 AMP_ENABLE, main.py:242-243), AdamW,
batch 256, AFF-Tiny-1/5 224 x 224 — beside this repo's own arm measured by bench.py on the same box.

    python tests/ref_cuda_e2e.py [--batch 256] [--steps 5] [--out profiles/r02_ref_cuda_e2e.json]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))
import torch
from oracle import aff_ops_ref as A, point_ref as P, ref_cuda as RC


class _RefOp(torch.autograd.Function):
    """one reference operator: forward / backward through the compiled reference module (clusten/clusten.py:17-92)"""

    @staticmethod
    def forward(ctx, op, a, b, idx):
        a, b, idx = a.contiguous(), b.contiguous(), idx.contiguous()
        if b.dtype != a.dtype:
            b = b.to(a.dtype)
        ctx.op = op
        ctx.save_for_backward(a, b, idx)
        return RC.forward(op, a, b, idx)

    @staticmethod
    def backward(ctx, go):
        a, b, idx = ctx.saved_tensors
        ga, gb = RC.backward(ctx.op, go, a, b, idx)
        return None, ga, gb, None


def _on(dev, fn, *a, **kw):
    with torch.device(dev):      # the oracle's tensor factories follow the operands onto the GPU
        return fn(*a, **kw)


class RefCudaOps(A.OracleOps):
    fused_attention = None

    @staticmethod
    def qk(q, k, idx):
        return _RefOp.apply("qk", q, k, idx)

    @staticmethod
    def av(attn, v, idx):
        return _RefOp.apply("av", attn, v, idx)

    @staticmethod
    def wf(w, f, idx):
        return _RefOp.apply("wf", w, f, idx)

    @staticmethod
    def space_filling_cluster(pos, m, h, w, no_reorder=False):
        return _on(pos.device, P.sfc_ref, pos, m, h, w)

    @staticmethod
    def build_nbhd(pos, mean, m, nbhd_size, rel_pos_width, clamp=False):
        return _on(pos.device, P.nbhd_ref, pos, mean, m, nbhd_size, rel_pos_width, clamp)

    @staticmethod
    def select_merge_centers(pos, learned_prob, stride, alpha, ds_rate, reserve_num, reserve_on=True):
        return _on(pos.device, P.select_ref, pos, learned_prob, stride, alpha, ds_rate, reserve_num, reserve_on)


def _knn_cuda(query, database, k, return_dist=False):
    from affb200.point_utils import knn_keops
    return knn_keops(query.float().contiguous(), database.float().contiguous(), k, return_dist=return_dist)


def ev(fn, steps, warm):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--model", default="aff_tiny_1_5th")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    if not RC.available():
        print(json.dumps({"unavailable": "oracle/_ref is not built"}))
        return
    P.knn_ref = _knn_cuda          # pykeops stand-in (see the module docstring)
    from affb200.aff_transformer import build_aff
    torch.backends.cudnn.benchmark = True
    torch.manual_seed(0)
    dev = torch.device("cuda:0")
    model = build_aff(args.model, ops=RefCudaOps()).to(dev).train()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=0.05)
    scaler = torch.amp.GradScaler("cuda")
    g = torch.Generator().manual_seed(1)
    x = torch.randn(args.batch, 3, 224, 224, generator=g).to(dev)
    y = torch.randint(0, 1000, (args.batch,), generator=g).to(dev)

    def train_step():
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.float16):
            loss = torch.nn.functional.cross_entropy(model(x).float(), y)
        scaler.scale(loss).backward()
        scaler.step(opt)
        scaler.update()

    t0 = time.time()
    ms_train = ev(train_step, args.steps, args.warmup)
    model.eval()

    def infer():
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.float16):
            model(x)
    ms_inf = ev(infer, args.steps, args.warmup)
    res = {"what": "reference CLUSTEN CUDA kernels (oracle/_ref) + stock torch glue, eager, fp16 autocast + GradScaler",
           "model": args.model, "batch": args.batch, "steps": args.steps,
           "train_ms_per_step": ms_train, "train_img_s": args.batch / ms_train * 1e3,
           "infer_ms_per_step": ms_inf, "infer_img_s": args.batch / ms_inf * 1e3,
           "knn": "affb200_knn2d in place of pykeops (absent)", "torch": torch.__version__,
           "gpu": torch.cuda.get_device_name(0), "wall_s": time.time() - t0}
    line = json.dumps(res)
    print(line)
    if args.out:
        with open(args.out, "w") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    main()
