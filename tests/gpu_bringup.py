"""GPU bring-up diagnostics (not a pytest file): runs a ladder of forward/backward cases through the public API,
each in its own subprocess (a device-side trap kills the CUDA context), and prints max errors against the oracle.

    python tests/gpu_bringup.py            # whole ladder
    python tests/gpu_bringup.py --case 3   # one case, in-process
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hazyresearch_flash-attention_b200"))

# (name, impl, b, h, sq, sk, d, dtype, causal, dropout, varlen, bwd)
CASES = [
    ("simt_small", 2, 2, 2, 96, 96, 64, "fp16", False, 0.0, True, True),
    ("tc_s128_d64", 0, 1, 1, 128, 128, 64, "fp16", False, 0.0, False, False),
    ("tc_s128_d128", 0, 1, 1, 128, 128, 128, "bf16", False, 0.0, False, False),
    ("tc_s256_d64", 0, 1, 2, 256, 256, 64, "fp16", False, 0.0, False, False),
    ("tc_s512_d128", 0, 2, 2, 512, 512, 128, "bf16", False, 0.0, False, False),
    ("tc_s1024_d64_causal", 0, 2, 2, 1024, 1024, 64, "fp16", True, 0.0, False, False),
    ("tc_s1024_d128_causal", 0, 2, 2, 1024, 1024, 128, "bf16", True, 0.0, False, False),
    ("tc_varlen_d64", 0, 4, 2, 200, 200, 64, "fp16", False, 0.0, True, False),
    ("tc_varlen_d80_causal", 0, 4, 2, 333, 333, 80, "fp16", True, 0.0, True, False),
    ("tc_cross_d96", 0, 3, 2, 257, 140, 96, "bf16", False, 0.0, True, False),
    ("tc_dropout_d64", 0, 2, 2, 384, 384, 64, "fp16", False, 0.17, True, False),
    ("tc_dropout_d128_causal", 0, 2, 2, 384, 384, 128, "bf16", True, 0.17, True, False),
    ("tc_d32", 0, 2, 2, 130, 130, 32, "fp16", False, 0.0, True, False),
    ("tc_bwd_d64", 0, 2, 2, 256, 256, 64, "fp16", True, 0.1, True, True),
    ("tc_bwd_d128", 0, 2, 2, 384, 384, 128, "bf16", False, 0.0, True, True),
    # backward ladder (indices 15..)
    ("bwd_1tile_d64", 0, 1, 1, 128, 128, 64, "fp16", False, 0.0, False, True),
    ("bwd_1tile_d128", 0, 1, 1, 128, 128, 128, "bf16", False, 0.0, False, True),
    ("bwd_2x2_d64", 0, 1, 2, 256, 256, 64, "fp16", False, 0.0, False, True),
    ("bwd_4x4_d128", 0, 2, 2, 512, 512, 128, "bf16", False, 0.0, False, True),
    ("bwd_causal_d128", 0, 2, 2, 512, 512, 128, "bf16", True, 0.0, False, True),
    ("bwd_varlen_d64", 0, 3, 2, 300, 300, 64, "fp16", False, 0.0, True, True),
    ("bwd_cross_causal_d96", 0, 3, 2, 257, 400, 96, "bf16", True, 0.0, True, True),
    ("bwd_dropout_d64", 0, 2, 2, 384, 384, 64, "fp16", False, 0.17, False, True),
    ("bwd_dropout_causal_varlen_d128", 0, 2, 2, 500, 500, 128, "bf16", True, 0.17, True, True),
    ("bwd_d32_long", 0, 1, 2, 2048, 2048, 32, "fp16", True, 0.0, False, True),
]


def run_case(idx):
    import torch
    import flash_attn_cuda
    from flash_attn.flash_attn_interface import flash_attn_unpadded_func
    from oracle import attention_ref as orc

    name, impl, b, h, sq, sk, d, dt, causal, p, varlen, bwd = CASES[idx]
    dtype = torch.float16 if dt == "fp16" else torch.bfloat16
    dev = "cuda"
    torch.manual_seed(idx)
    flash_attn_cuda._impl = impl
    q = torch.randn(b, sq, h, d, device=dev, dtype=dtype)
    k = torch.randn(b, sk, h, d, device=dev, dtype=dtype)
    v = torch.randn(b, sk, h, d, device=dev, dtype=dtype)
    if varlen:
        qmask = orc.generate_random_padding_mask(sq, b, dev, mode="third")
        kmask = qmask if sq == sk else orc.generate_random_padding_mask(sk, b, dev, mode="third")
    else:
        qmask = torch.ones(b, sq, dtype=torch.bool, device=dev)
        kmask = torch.ones(b, sk, dtype=torch.bool, device=dev)
    ql, kl = qmask.sum(-1), kmask.sum(-1)
    cu_q = torch.nn.functional.pad(ql.cumsum(0), (1, 0)).int()
    cu_k = torch.nn.functional.pad(kl.cumsum(0), (1, 0)).int()
    qu = q[qmask].clone().requires_grad_(bwd)
    ku = k[kmask].clone().requires_grad_(bwd)
    vu = v[kmask].clone().requires_grad_(bwd)
    t0 = time.time()
    out_u, lse, S = flash_attn_unpadded_func(qu, ku, vu, cu_q, cu_k, sq, sk, p, causal=causal,
                                             return_attn_probs=True)
    torch.cuda.synchronize()
    res = {"case": name, "fwd_s": round(time.time() - t0, 3)}
    out = torch.zeros(b, sq, h, d, device=dev, dtype=dtype)
    out[qmask] = out_u.detach()
    S = S[:, :, :sq, :sk]
    dmask = ~torch.signbit(S.float())
    qf, kf, vf = [x.clone().float().requires_grad_(bwd) for x in (q, k, v)]
    out_ref, attn_ref = orc.attention_ref(qf, kf, vf, qmask, kmask, p, dmask, causal=causal)
    out_pt, _ = orc.attention_ref(q, k, v, qmask, kmask, p, dmask, causal=causal, upcast=False, reorder_ops=True)
    lse_ref = orc.softmax_lse_ref(q, k, qmask, kmask, causal)
    valid = qmask[:, None, :].expand(b, h, sq)
    lse_c = lse[:, :, :sq]
    finite = valid & torch.isfinite(lse_ref)
    res["lse_err"] = (lse_c - lse_ref)[finite].abs().max().item() if finite.any() else 0.0
    res["o_err"] = (out.float() - out_ref.float()).abs().max().item()
    res["o_pt_err"] = (out_pt.float() - out_ref.float()).abs().max().item()
    res["attn_err"] = (S.abs().float() - attn_ref.float()).abs().max().item()
    res["o_nan"] = bool(torch.isnan(out).any().item())
    if p > 0:
        res["drop_frac"] = orc.get_dropout_fraction(dmask, qmask, kmask, causal).item()
    if bwd:
        g = torch.randn_like(out_u)
        dqu, dku, dvu = torch.autograd.grad(out_u, (qu, ku, vu), g)
        gp = torch.zeros(b, sq, h, d, device=dev, dtype=dtype)
        gp[qmask] = g
        dq_ref, dk_ref, dv_ref = torch.autograd.grad(out_ref, (qf, kf, vf), gp.float())
        qp, kp, vp = [x.clone().requires_grad_() for x in (q, k, v)]
        out_pt2, _ = orc.attention_ref(qp, kp, vp, qmask, kmask, p, dmask, causal=causal, upcast=False, reorder_ops=True)
        pt_grads = torch.autograd.grad(out_pt2, (qp, kp, vp), gp)
        bwd_ok = True
        for nm, got, ref, ptg, m in (("dq", dqu, dq_ref, pt_grads[0], qmask), ("dk", dku, dk_ref, pt_grads[1], kmask),
                                     ("dv", dvu, dv_ref, pt_grads[2], kmask)):
            res[nm + "_err"] = (got.float() - ref[m]).abs().max().item()
            res[nm + "_pt_err"] = (ptg.float()[m] - ref[m]).abs().max().item()
            res[nm + "_max"] = ref[m].abs().max().item()
            res[nm + "_nan"] = bool(torch.isnan(got).any().item())
            bwd_ok = bwd_ok and res[nm + "_err"] <= 2 * res[nm + "_pt_err"] + 1e-5 and not res[nm + "_nan"]
            if not (res[nm + "_err"] <= 2 * res[nm + "_pt_err"] + 1e-5):
                full = torch.zeros_like(ref)
                full[m] = got.float()
                e = (full - ref * m[..., None, None]).abs()
                print(nm, "err per 64-row block (b0,h0):", e[0, :, 0].amax(-1).reshape(-1)[: (e.shape[1] // 64) * 64].reshape(-1, 64).amax(-1))
                print(nm, "err per 8-col block (b0,h0):", e[0, :, 0].amax(0).reshape(-1, 8).amax(-1))
                print(nm, "got[0,:4,0,:8]\n", full[0, :4, 0, :8], "\nref\n", ref[0, :4, 0, :8])
        res["bwd_ok"] = bwd_ok
    res["ok"] = (res["o_err"] <= 2 * res["o_pt_err"] + 1e-5) and res["lse_err"] < 2e-3 and not res["o_nan"] \
        and res.get("bwd_ok", True)
    print("RESULT " + json.dumps(res), flush=True)
    if not res["ok"]:
        torch.set_printoptions(precision=4, linewidth=200)
        print("out[0,:4,0,:8]\n", out[0, :4, 0, :8].float(), "\nref\n", out_ref[0, :4, 0, :8].float())
        print("lse[0,0,:8]", lse_c[0, 0, :8], "ref", lse_ref[0, 0, :8])
        err = (out.float() - out_ref.float()).abs()
        print("err per row-block of 32 (b0,h0):", err[0, :, 0].amax(-1).reshape(-1)[: (sq // 32) * 32].reshape(-1, 32).amax(-1))
        print("err per col-block of 8 (b0,h0):", err[0, :, 0].amax(0).reshape(-1, 8).amax(-1))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", type=int, default=None)
    ap.add_argument("--only", type=str, default=None, help="comma separated case indices")
    args = ap.parse_args()
    if args.case is not None:
        run_case(args.case)
        return
    idxs = range(len(CASES)) if args.only is None else [int(x) for x in args.only.split(",")]
    for i in idxs:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--case", str(i)], capture_output=True,
                               text=True, timeout=180)
            tail = "\n".join((r.stdout + "\n" + r.stderr).strip().splitlines()[-30:])
            status = "exit=%d" % r.returncode
        except subprocess.TimeoutExpired as e:
            tail, status = str(e), "TIMEOUT"
        print(f"=== case {i} {CASES[i][0]} [{status}, {time.time() - t0:.1f}s]\n{tail}\n", flush=True)


if __name__ == "__main__":
    main()
