"""Generates the golden fixtures under tests/golden/ by running the REFERENCE's own Python functions
(read from /root/reference at generation time, never copied into this repo) on seeded CPU inputs.

    python tests/golden/make_golden.py           # needs /root/reference; writes tests/golden/*.pt

What is pinned:
  attention_ref / attention_kvpacked_ref / attention_qkvpacked_ref   tests/test_flash_attn.py:115-173
  attention_blocksparse_ref                                          tests/test_flash_attn.py:189-215
  generate_random_padding_mask, get_dropout_fraction                 tests/test_flash_attn.py:17-26, 300-329
  convert_blockmask                                                  flash_attn/flash_blocksparse_attn_interface.py:8-40
  unpad_input / pad_input                                            flash_attn/bert_padding.py:99-134
  RotaryEmbedding / RotaryEmbedding2D                                flash_attn/rotary.py:45-135
The reference holds no golden vectors of its own, so these are outputs of the reference code
itself.  Gradients are autograd of the reference's attention_ref (what its commented-out asserts compare with).
"""
import ast
import math
import os
import sys
import types

import torch
import torch.nn.functional as F
from einops import rearrange, repeat

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def load_functions(path, names, extra_ns=None):
    """exec only the named top-level functions/classes of a reference file (its module-level code needs a GPU
    and the reference's native extension)."""
    tree = ast.parse(open(path).read())
    keep = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in names]
    ns = {"math": math, "torch": torch, "F": F, "rearrange": rearrange, "repeat": repeat, "nn": torch.nn}
    ns.update(extra_ns or {})
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    return ns


def pack_bool(m):
    """bool tensor -> (shape, packed uint8 bits): keeps the fixtures small."""
    import numpy as np
    return dict(shape=tuple(m.shape), bits=torch.from_numpy(np.packbits(m.numpy().reshape(-1))))


def small_or_none(x, limit=120_000):
    return x if x.numel() <= limit else None


def main():
    t = load_functions(f"{REF}/tests/test_flash_attn.py",
                       {"generate_random_padding_mask", "attention_ref", "attention_kvpacked_ref",
                        "attention_qkvpacked_ref", "attention_blocksparse_ref", "get_dropout_fraction"})
    bs = load_functions(f"{REF}/flash_attn/flash_blocksparse_attn_interface.py", {"convert_blockmask"})
    import numpy as np
    bp = load_functions(f"{REF}/flash_attn/bert_padding.py",
                        {"IndexFirstAxis", "IndexPutFirstAxis", "unpad_input", "pad_input"}, {"np": np})
    bp["index_first_axis"] = bp["IndexFirstAxis"].apply
    bp["index_put_first_axis"] = bp["IndexPutFirstAxis"].apply

    cases = []
    g = torch.Generator().manual_seed(1234)
    # (batch, seqlen_q, seqlen_k, heads, d, causal, dropout_p, dtype, kind)
    specs = [
        (2, 97, 97, 2, 32, False, 0.0, torch.float16, "qkv"),
        (2, 128, 128, 2, 64, True, 0.17, torch.float16, "qkv"),
        (2, 200, 200, 1, 56, True, 0.0, torch.bfloat16, "qkv"),
        (1, 257, 257, 2, 128, False, 0.17, torch.bfloat16, "qkv"),
        (2, 113, 203, 1, 80, False, 0.0, torch.float16, "kv"),
        (2, 203, 113, 1, 96, True, 0.17, torch.float16, "kv"),
        (3, 64, 300, 1, 64, True, 0.0, torch.bfloat16, "sep"),
        (1, 384, 384, 1, 128, True, 0.0, torch.float16, "sep"),
    ]
    for (b, sq, sk, h, d, causal, p, dtype, kind) in specs:
        torch.manual_seed(b * 1000 + sq + sk + d)
        q = torch.randn(b, sq, h, d, generator=g).to(dtype)
        k = torch.randn(b, sk, h, d, generator=g).to(dtype)
        v = torch.randn(b, sk, h, d, generator=g).to(dtype)
        if kind == "qkv":
            qmask = t["generate_random_padding_mask"](sq, b, "cpu", mode="random")
            kmask = qmask
        else:
            qmask = t["generate_random_padding_mask"](sq, b, "cpu", mode="random")
            kmask = t["generate_random_padding_mask"](sk, b, "cpu", mode="third")
        dmask = (torch.rand(b, h, sq, sk, generator=g) >= p) if p > 0 else torch.ones(b, h, sq, sk, dtype=torch.bool)
        qf, kf, vf = [x.clone().requires_grad_() for x in (q, k, v)]
        out_ref, attn_ref = t["attention_ref"](qf, kf, vf, qmask, kmask, p, dmask, causal=causal)
        qp, kp, vp = [x.clone().requires_grad_() for x in (q, k, v)]
        out_pt, attn_pt = t["attention_ref"](qp, kp, vp, qmask, kmask, p, dmask, causal=causal, upcast=False,
                                             reorder_ops=True)
        dout = torch.randn(out_ref.shape, generator=g).to(dtype)
        dq, dk, dv = torch.autograd.grad(out_ref, (qf, kf, vf), dout)
        # the 16-bit PyTorch path's own gradient error: the denominator of the reference's 2x rule (:407-418)
        grads_pt = torch.autograd.grad(out_pt, (qp, kp, vp), dout)
        grad_pt_maxerr = [(a.float() - r.float()).abs().max().item() for a, r in zip(grads_pt, (dq, dk, dv))]
        frac = t["get_dropout_fraction"](dmask.clone(), qmask, kmask, causal=causal).item()
        cases.append(dict(b=b, sq=sq, sk=sk, h=h, d=d, causal=causal, p=p, dtype=str(dtype), kind=kind,
                          q=q, k=k, v=v, qmask=qmask, kmask=kmask, dmask=pack_bool(dmask), dout=dout,
                          out_ref=out_ref.detach(), attn_ref=small_or_none(attn_ref.detach()),
                          out_pt_maxerr=(out_pt.float() - out_ref.detach().float()).abs().max().item(),
                          attn_pt_maxerr=(attn_pt.float() - attn_ref.detach().float()).abs().max().item(),
                          dq=dq, dk=dk, dv=dv, grad_pt_maxerr=grad_pt_maxerr, dropout_fraction=frac))
    torch.save(cases, os.path.join(OUT, "attention_ref_cases.pt"))

    # blocksparse oracle + blockmask conversion
    bcases = []
    for (b, s, h, d, sparsity, p) in [(2, 256, 2, 32, 0.5, 0.0), (1, 512, 1, 64, 0.4, 0.17), (1, 768, 1, 64, 0.7, 0.0)]:
        qkv = torch.randn(b, s, 3, h, d, generator=g).to(torch.float16)
        blockmask = torch.rand(s // 16, s // 256, generator=g) < sparsity
        blockmask[:, 0] |= ~blockmask.any(dim=1)   # every row attends somewhere (no NaN rows in the reference math)
        amask = t["generate_random_padding_mask"](s, b, "cpu", mode="random")
        dmask = (torch.rand(b, h, s, s, generator=g) >= p) if p > 0 else torch.ones(b, h, s, s, dtype=torch.bool)
        out, attn = t["attention_blocksparse_ref"](qkv, blockmask, amask, p, dmask)
        out32, _ = t["attention_blocksparse_ref"](qkv.float(), blockmask, amask, p, dmask)   # same math, fp32 result
        bcases.append(dict(out32=out32, b=b, s=s, h=h, d=d, p=p, qkv=qkv, blockmask=blockmask, amask=amask, dmask=pack_bool(dmask),
                           out=out, attn=small_or_none(attn, 300_000), converted=bs["convert_blockmask"](blockmask.clone(), False)))
    conv = []
    for (nrow, ncol, sp) in [(8, 1, 0.5), (16, 2, 0.3), (64, 4, 0.3), (128, 8, 0.7), (33, 5, 0.2), (16, 3, 1.0), (16, 3, 0.0)]:
        bm = torch.rand(nrow, ncol, generator=g) < sp
        conv.append(dict(blockmask=bm, converted=bs["convert_blockmask"](bm.clone(), False)))
    torch.save(dict(attn=bcases, convert=conv), os.path.join(OUT, "blocksparse_cases.pt"))

    # bert_padding
    pcases = []
    for (b, s, w) in [(3, 17, 8), (4, 64, 24)]:
        x = torch.randn(b, s, w, generator=g)
        mask = t["generate_random_padding_mask"](s, b, "cpu", mode="third")
        xu, idx, cu, mx = bp["unpad_input"](x, mask)
        back = bp["pad_input"](xu, idx, b, s)
        pcases.append(dict(x=x, mask=mask, x_unpad=xu, indices=idx, cu_seqlens=cu, max_seqlen=mx, repadded=back))
    torch.save(pcases, os.path.join(OUT, "bert_padding_cases.pt"))

    # rotary (the reference file scripts a function with torch.jit; load the whole module under a private name)
    import importlib.util
    spec = importlib.util.spec_from_file_location("_ref_rotary", f"{REF}/flash_attn/rotary.py")
    rot = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(rot)
    q = torch.randn(2, 16, 3, 32, generator=g)
    k = torch.randn(2, 16, 3, 32, generator=g)
    q1, k1 = rot.RotaryEmbedding(32)(q, k, seq_dimension=-3)
    q2, k2 = rot.RotaryEmbedding2D(32)(q, k, seq_dimension=-3)
    torch.save(dict(q=q, k=k, q1=q1, k1=k1, q2=q2, k2=k2), os.path.join(OUT, "rotary_cases.pt"))
    print("wrote", sorted(f for f in os.listdir(OUT) if f.endswith(".pt")))


if __name__ == "__main__":
    main()
