"""ncu target: a few launches of the attention core at one geometry.   python scripts/prof_attn.py [L14|B16|text] [impl]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
lib, check, ptr, stream = pkg._lib.lib, pkg._lib.check, pkg._lib.ptr, pkg._lib.current_stream_ptr
dev = torch.device("cuda:0")
geo = {"L14": (128, 257, 1024, 16, 0), "B16": (256, 197, 768, 12, 0), "text": (64, 77, 512, 8, 1),
       "B32": (256, 50, 768, 12, 0)}[sys.argv[1] if len(sys.argv) > 1 else "L14"]
impl = int(sys.argv[2]) if len(sys.argv) > 2 else 2
B, L, d, heads, causal = geo
qkv = torch.randn((B * L, 3 * d), device=dev).to(torch.bfloat16)
out = torch.empty((B * L, d), dtype=torch.bfloat16, device=dev)
for _ in range(4):
    check(lib.evr_attention(ptr(qkv), ptr(out), B, L, d, heads, causal, impl, stream()), "attn")
torch.cuda.synchronize()
print("ok")
