"""Second baseline: stock PyTorch eager on the same B200 for the two bench workloads.

Not the product and not the oracle: a plain `torch.nn` ViT (cuBLAS GEMMs, SDPA attention, ATen LayerNorm) in bf16 with
random weights, and `torch.matmul` + `torch.topk` for the scan, i.e. what the reference would run if its tensors were
simply moved to `cuda` (`embedding_service.py:41, 314-320` with torch instead of numpy).  Prints one JSON line per
workload; the numbers are quoted in DESIGN.md next to ours.  Timing: CUDA events, 3 warm-up + `--steps` timed steps.
"""
import argparse
import json

import torch
import torch.nn as nn
import torch.nn.functional as F


class Block(nn.Module):
    def __init__(self, d, h):
        super().__init__()
        self.h = h
        self.ln_1, self.ln_2 = nn.LayerNorm(d), nn.LayerNorm(d)
        self.qkv, self.out = nn.Linear(d, 3 * d), nn.Linear(d, d)
        self.fc, self.proj = nn.Linear(d, 4 * d), nn.Linear(4 * d, d)

    def forward(self, x):
        B, L, d = x.shape
        q, k, v = self.qkv(self.ln_1(x)).view(B, L, 3, self.h, d // self.h).permute(2, 0, 3, 1, 4)
        a = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(B, L, d)
        x = x + self.out(a)
        y = self.fc(self.ln_2(x))
        return x + self.proj(y * torch.sigmoid(1.702 * y))


class EagerViT(nn.Module):
    def __init__(self, patch=32, d=768, layers=12, heads=12, out=512, res=224):
        super().__init__()
        self.conv = nn.Conv2d(3, d, patch, patch, bias=False)
        self.cls = nn.Parameter(torch.randn(d) * d ** -0.5)
        self.pos = nn.Parameter(torch.randn((res // patch) ** 2 + 1, d) * d ** -0.5)
        self.ln_pre, self.ln_post = nn.LayerNorm(d), nn.LayerNorm(d)
        self.blocks = nn.Sequential(*[Block(d, heads) for _ in range(layers)])
        self.proj = nn.Parameter(torch.randn(d, out) * d ** -0.5)
        self.register_buffer("mean", torch.tensor([0.48145466, 0.4578275, 0.40821073]).view(1, 3, 1, 1) * 255)
        self.register_buffer("std", torch.tensor([0.26862954, 0.26130258, 0.27577711]).view(1, 3, 1, 1) * 255)

    @torch.no_grad()
    def load_openai_state_dict(self, sd, prefix="visual."):
        """Same weights as the product's tower: openai/CLIP VisionTransformer keys -> this module's names."""
        def put(param, key):
            param.copy_(sd[prefix + key].reshape(param.shape))
        put(self.conv.weight, "conv1.weight")
        put(self.cls, "class_embedding")
        put(self.pos, "positional_embedding")
        put(self.proj, "proj")
        for ln, key in ((self.ln_pre, "ln_pre"), (self.ln_post, "ln_post")):
            put(ln.weight, key + ".weight")
            put(ln.bias, key + ".bias")
        for i, blk in enumerate(self.blocks):
            b = f"transformer.resblocks.{i}."
            for mod, key in ((blk.ln_1, "ln_1"), (blk.ln_2, "ln_2"), (blk.out, "attn.out_proj"), (blk.fc, "mlp.c_fc"),
                             (blk.proj, "mlp.c_proj")):
                put(mod.weight, b + key + ".weight")
                put(mod.bias, b + key + ".bias")
            put(blk.qkv.weight, b + "attn.in_proj_weight")
            put(blk.qkv.bias, b + "attn.in_proj_bias")
        return self

    def forward(self, frames_u8):  # [B, H, W, 3] uint8, as bench.py feeds our tower
        x = ((frames_u8.permute(0, 3, 1, 2).to(self.mean.dtype) - self.mean) / self.std)
        x = self.conv(x).flatten(2).transpose(1, 2)
        x = torch.cat([self.cls.expand(x.shape[0], 1, -1), x], 1) + self.pos
        x = self.blocks(self.ln_pre(x))
        e = self.ln_post(x[:, 0]) @ self.proj
        e = F.normalize(e.float(), dim=-1)
        return e if self.proj.dtype == torch.float32 else e.half()


def timed(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


@torch.no_grad()
def encode(args):
    dev = torch.device("cuda:0")
    geo = dict(patch=32, d=768, layers=12, heads=12, out=512) if args.model == "ViT-B/32" else \
        dict(patch=14, d=1024, layers=24, heads=16, out=768)
    m = EagerViT(**geo).to(dev).to(torch.bfloat16).eval()
    gen = torch.Generator(device=dev).manual_seed(0)
    # 8 resident batches (308 MB > L2) rotated, like bench.py's value leg
    pool = [torch.randint(0, 256, (args.batch, 224, 224, 3), dtype=torch.uint8, device=dev, generator=gen)
            for _ in range(8)]
    i = [0]

    def step():
        m(pool[i[0] % 8])
        i[0] += 1
    ms = timed(step, args.steps)
    return {"impl": "torch_eager", "metric": "clip_image_encode_throughput", "value": args.batch / ms * 1e3,
            "unit": "frames/s", "ms_per_step": ms, "dtype": "bf16",
            "config": {"workload": f"{args.model} B={args.batch} 224x224 uint8 resident", "torch": torch.__version__}}


@torch.no_grad()
def search(args):
    dev = torch.device("cuda:0")
    N, D, Q, k, chunk = args.rows, 512, args.queries, 100, args.chunk
    gen = torch.Generator(device=dev).manual_seed(1)
    store = torch.empty((N, D), dtype=torch.float16, device=dev)
    for lo in range(0, N, 1 << 20):
        blk = torch.randn((min(1 << 20, N - lo), D), device=dev, generator=gen)
        store[lo:lo + blk.shape[0]] = F.normalize(blk, dim=-1).half()
    q = F.normalize(torch.randn((Q, D), device=dev, generator=gen), dim=-1).half()

    def step():
        best_s = best_i = None
        for lo in range(0, N, chunk):
            s = q @ store[lo:lo + chunk].T                    # [Q, chunk] fp16 (cuBLAS, fp32 accumulate)
            ts, ti = torch.topk(s, k, dim=1)
            ti += lo
            if best_s is None:
                best_s, best_i = ts, ti
            else:
                cs, ci = torch.cat([best_s, ts], 1), torch.cat([best_i, ti], 1)
                best_s, sel = torch.topk(cs, k, dim=1)
                best_i = torch.gather(ci, 1, sel)
        return best_s, best_i
    ms = timed(step, args.steps, warmup=2)
    return {"impl": "torch_eager", "metric": "topk_search_throughput", "value": Q / ms * 1e3, "unit": "queries/s",
            "ms_per_step": ms, "dtype": "f16",
            "config": {"workload": f"Q={Q} N={N} D={D} k={k} matmul+topk, {chunk}-row chunks",
                       "torch": torch.__version__}}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="encode,search")
    ap.add_argument("--model", default="ViT-B/32")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--queries", type=int, default=10_000)
    ap.add_argument("--chunk", type=int, default=500_000)
    a = ap.parse_args()
    for w in a.what.split(","):
        if w == "search":
            a.steps = min(a.steps, 3)
        print(json.dumps({"encode": encode, "search": search}[w](a)), flush=True)
        torch.cuda.empty_cache()
