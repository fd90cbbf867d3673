#!/usr/bin/env python
"""Per-kernel roofline bench on one B200: the tcgen05 GEMM on the Llama-3.1-8B training shapes (all
three operand layouts, every tile configuration) against cuBLAS (torch.matmul), the attention kernels
against flash-attn 2, and the memory-bound kernels against the measured copy bandwidth.

CUDA-event timing, >= 3 warm-ups, inputs rotated through a set larger than L2 between iterations.
Writes ``gpurun_out/kernel_bench.json`` and prints a table.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from b200train import ops  # noqa: E402
from b200train.ops.linear import gemm  # noqa: E402

PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}   # the profiling recipe's fallback; replaced by the measured file
try:
    PEAKS.update(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))))
except Exception:
    pass


def timeit(fn, n_rot, iters=10, warm=3):
    for i in range(warm):
        fn(i % n_rot)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i % n_rot)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def rnd(*shape):
    return (torch.randn(*shape, device="cuda") * 0.1).to(torch.bfloat16)


def bench_gemm(results, quick=False):
    T = 4096
    # (name, M, N, K, layout) -- layouts: fwd kk, dgrad k/mn, wgrad mn/mn
    shapes = [
        ("qkv_fwd", T, 6144, 4096, "kk"), ("o_fwd", T, 4096, 4096, "kk"),
        ("gateup_fwd", T, 28672, 4096, "kk"), ("down_fwd", T, 4096, 14336, "kk"),
        ("head_fwd", T, 128256, 4096, "kk"),
        ("qkv_dgrad", T, 4096, 6144, "kmn"), ("o_dgrad", T, 4096, 4096, "kmn"),
        ("gateup_dgrad", T, 4096, 28672, "kmn"),
        ("down_dgrad", T, 14336, 4096, "kmn"),
        ("qkv_wgrad", 6144, 4096, T, "mnmn"), ("o_wgrad", 4096, 4096, T, "mnmn"),
        ("gateup_wgrad", 28672, 4096, T, "mnmn"),
        ("down_wgrad", 4096, 14336, T, "mnmn"), ("head_wgrad", 128256, 4096, T, "mnmn"),
    ]
    if quick:
        shapes = shapes[:4]
    configs = [("c1n256", 1 | (256 << 8)), ("c2n256", 2 | (256 << 8)), ("c1n128", 1 | (128 << 8)),
               ("c2n128", 2 | (128 << 8)), ("c2n256_whole_tail", 2 | (256 << 8) | (1 << 24))]
    for name, M, N, K, layout in shapes:
        a_mn = layout in ("mnmn",)
        b_mn = layout in ("kmn", "mnmn")
        bytes_ = 2 * (M * K + N * K + M * N)
        n_rot = max(2, int(300e6 // bytes_) + 1)
        A = [rnd(K, M) if a_mn else rnd(M, K) for _ in range(n_rot)]
        Bm = [rnd(K, N) if b_mn else rnd(N, K) for _ in range(n_rot)]
        C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        flops = 2.0 * M * N * K
        row = {"shape": [M, N, K], "layout": layout}

        def cublas(i):
            a = A[i].t() if a_mn else A[i]
            b = Bm[i] if b_mn else Bm[i].t()
            torch.matmul(a, b, out=C)

        ms = timeit(cublas, n_rot)
        row["cublas_tflops"] = flops / ms / 1e9
        best = 0
        for cname, cfg in configs:
            try:
                ms = timeit(lambda i: gemm(A[i], Bm[i], a_mn=a_mn, b_mn=b_mn, out=C, config=cfg), n_rot)
                row[cname + "_tflops"] = flops / ms / 1e9
                best = max(best, row[cname + "_tflops"])
            except Exception as e:
                row[cname + "_tflops"] = f"ERR {e}"
        row["best_vs_cublas"] = best / row["cublas_tflops"]
        row["best_frac_of_measured_peak"] = best / PEAKS["bf16_tflops"]
        results["gemm"][name] = row
        print(name, {k: (round(v, 3) if isinstance(v, float) else v) for k, v in row.items()}, flush=True)
        del A, Bm, C


def bench_attention(results):
    from b200train.ops.attention import attn_bwd_raw, attn_fwd_raw

    for name, B, S, nq, nkv, D in [("llama8b_s4096", 1, 4096, 32, 8, 128), ("llama8b_s8192", 1, 8192, 32, 8, 128),
                                   ("gpt2_s1024", 16, 1024, 12, 12, 64)]:
        T = B * S
        qkv = rnd(T, (nq + 2 * nkv) * D) * 10
        cu = (torch.arange(0, B + 1, device="cuda", dtype=torch.int32) * S)
        q, k, v = qkv[:, :nq * D], qkv[:, nq * D:(nq + nkv) * D], qkv[:, (nq + nkv) * D:]
        flops_fwd = 4.0 * B * S * S * nq * D / 2  # causal
        row = {}
        ms = timeit(lambda i: attn_fwd_raw(q, k, v, cu, S, nq, nkv, D, D ** -0.5), 1)
        row["ours_fwd_ms"] = ms
        row["ours_fwd_tflops"] = flops_fwd / ms / 1e9
        try:
            from flash_attn import flash_attn_func

            q4 = q.reshape(B, S, nq, D).contiguous()
            k4 = k.reshape(B, S, nkv, D).contiguous()
            v4 = v.reshape(B, S, nkv, D).contiguous()
            ms = timeit(lambda i: flash_attn_func(q4, k4, v4, causal=True), 1)
            row["fa2_fwd_ms"] = ms
            row["fa2_fwd_tflops"] = flops_fwd / ms / 1e9
            q4.requires_grad_(), k4.requires_grad_(), v4.requires_grad_()
            o4 = flash_attn_func(q4, k4, v4, causal=True)
            do4 = torch.randn_like(o4)
            ms = timeit(lambda i: torch.autograd.grad(o4, (q4, k4, v4), do4, retain_graph=True), 1)
            row["fa2_bwd_ms"] = ms
            row["fa2_bwd_tflops"] = 2.5 * flops_fwd / ms / 1e9
        except Exception as e:
            row["fa2"] = f"ERR {e}"
        try:
            o, lse = attn_fwd_raw(q, k, v, cu, S, nq, nkv, D, D ** -0.5)
            do = torch.randn_like(o)
            dqkv = torch.empty_like(qkv)
            dq, dk, dv = dqkv[:, :nq * D], dqkv[:, nq * D:(nq + nkv) * D], dqkv[:, (nq + nkv) * D:]
            ms = timeit(lambda i: attn_bwd_raw(do, q, k, v, o, lse, dq, dk, dv, cu, S, nq, nkv, D, D ** -0.5), 1)
            row["ours_bwd_ms"] = ms
            row["ours_bwd_tflops"] = 2.5 * flops_fwd / ms / 1e9
        except Exception as e:
            row["ours_bwd"] = f"ERR {e}"
        results["attention"][name] = row
        print(name, {k: (round(v, 3) if isinstance(v, float) else v) for k, v in row.items()}, flush=True)


def bench_elementwise(results):
    T, H, I = 4096, 4096, 14336
    n_rot = 6
    xs = [rnd(T, H) for _ in range(n_rot)]
    rs = [rnd(T, H) for _ in range(n_rot)]
    w = rnd(H)
    ms = timeit(lambda i: ops.rmsnorm(xs[i], w, 1e-5, rs[i]), n_rot)
    results["elementwise"]["rmsnorm_res_fwd"] = {"ms": ms, "gbs": 4 * T * H * 2 / ms / 1e6}
    # raw C-ABI calls for the backward kernels (no autograd around them)
    from b200train import _C

    st = torch.cuda.current_stream().cuda_stream
    rstd = torch.rand(T, device="cuda") + 0.5
    dx = torch.empty(T, H, device="cuda", dtype=torch.bfloat16)
    dw = torch.empty(H, device="cuda", dtype=torch.bfloat16)
    nrows = _C.load().b2_rmsnorm_bwd_partial_rows(T)
    partial = torch.empty(nrows * H, device="cuda", dtype=torch.float32)
    yb, hb = torch.empty_like(xs[0]), torch.empty_like(xs[0])
    ms = timeit(lambda i: _C.call("b2_rmsnorm_fwd", yb.data_ptr(), rstd.data_ptr(), hb.data_ptr(), xs[i].data_ptr(),
                                  rs[i].data_ptr(), w.data_ptr(), T, H, 1e-5, st), n_rot)
    results["elementwise"]["rmsnorm_res_fwd_raw"] = {"ms": ms, "gbs": 4 * T * H * 2 / ms / 1e6}
    ms = timeit(lambda i: _C.call("b2_rmsnorm_bwd", dx.data_ptr(), dw.data_ptr(), partial.data_ptr(),
                                  xs[i].data_ptr(), rs[i].data_ptr(), w.data_ptr(), rstd.data_ptr(),
                                  rs[(i + 1) % n_rot].data_ptr(), T, H, 0, st), n_rot)
    results["elementwise"]["rmsnorm_res_bwd"] = {"ms": ms, "gbs": 4 * T * H * 2 / ms / 1e6}
    gus = [rnd(T, 2 * I) for _ in range(3)]
    ms = timeit(lambda i: ops.swiglu(gus[i]), 3)
    results["elementwise"]["swiglu_fwd"] = {"ms": ms, "gbs": 3 * T * I * 2 / ms / 1e6}
    douts = [rnd(T, I) for _ in range(3)]
    dgu = torch.empty(T, 2 * I, device="cuda", dtype=torch.bfloat16)
    ms = timeit(lambda i: _C.call("b2_swiglu_bwd", dgu.data_ptr(), douts[i].data_ptr(), gus[i].data_ptr(), T, I, st), 3)
    results["elementwise"]["swiglu_bwd"] = {"ms": ms, "gbs": 5 * T * I * 2 / ms / 1e6}
    del douts, dgu
    n = 218_112_000
    p, g, m, v = (rnd(n) for _ in range(4))
    ms = timeit(lambda i: ops.adamw_step(p, g, m, v, 1e-4, step=3), 1)
    results["elementwise"]["adamw_layer"] = {"ms": ms, "gbs": 7 * n * 2 / ms / 1e6}
    from b200train.ops.cross_entropy import linear_cross_entropy

    x = rnd(T, H).requires_grad_()
    wv = rnd(128256, H).requires_grad_()
    labels = torch.randint(0, 128256, (T,), device="cuda")

    def ce(i):
        loss = linear_cross_entropy(x, wv, labels)
        loss.backward()

    ms = timeit(ce, 1, iters=5)
    results["elementwise"]["head_ce_fwd_bwd"] = {"ms": ms, "tflops": 6.0 * T * H * 128256 / ms / 1e9}
    for k, r in results["elementwise"].items():
        if "gbs" in r:
            r["frac_of_measured_copy_bw"] = r["gbs"] / PEAKS["hbm_gbs"]
        print(k, {a: round(b, 3) for a, b in r.items()}, flush=True)


def bench_fused(results):
    """GEMMs with fused epilogues against the kernel sequences they replace (one Llama-3.1-8B layer)."""
    from b200train import _C
    from b200train.ops.dispatch import stream

    T, H, I = 4096, 4096, 14336
    n_rot = 2
    xs = [rnd(T, H) for _ in range(n_rot)]
    w_gu, w_d, w_qkv = rnd(2 * I, H), rnd(H, I), rnd(6144, H)
    gu = torch.empty(T, 2 * I, device="cuda", dtype=torch.bfloat16)
    h = torch.empty(T, I, device="cuda", dtype=torch.bfloat16)
    nul = (None, 0, 0, 0, 0, 0, 0, 0)
    st = stream(gu)
    r = results["fused"]
    flops = 2.0 * T * 2 * I * H
    ms = timeit(lambda i: _C.call("b2_gemm_swiglu", gu.data_ptr(), h.data_ptr(), xs[i].data_ptr(), w_gu.data_ptr(), T, I,
                                  H, H, H, 2 * I, I, *nul, st), n_rot)
    r["gateup_swiglu_fused"] = {"ms": ms, "tflops": flops / ms / 1e9}
    ms = timeit(lambda i: ops.swiglu(gemm(xs[i], w_gu, out=gu)), n_rot)
    r["gateup_then_swiglu"] = {"ms": ms, "tflops": flops / ms / 1e9}
    dys = [rnd(T, H) for _ in range(n_rot)]
    dgu = torch.empty_like(gu)
    flops = 2.0 * T * I * H
    ms = timeit(lambda i: _C.call("b2_gemm_dswiglu", dgu.data_ptr(), dys[i].data_ptr(), w_d.data_ptr(), gu.data_ptr(), T,
                                  I, H, H, I, 2 * I, 2 * I, *nul, st), n_rot)
    r["down_dgrad_dswiglu_fused"] = {"ms": ms, "tflops": flops / ms / 1e9}
    dh = torch.empty(T, I, device="cuda", dtype=torch.bfloat16)

    def unfused_bwd(i):
        gemm(dys[i], w_d, b_mn=True, out=dh)
        _C.call("b2_swiglu_bwd", dgu.data_ptr(), dh.data_ptr(), gu.data_ptr(), T, I, st)

    ms = timeit(unfused_bwd, n_rot)
    r["down_dgrad_then_swiglu_bwd"] = {"ms": ms, "tflops": flops / ms / 1e9}
    cos, sin = ops.rope_tables(128, 8192, 500000.0, None, device="cuda")
    qkv = torch.empty(T, 6144, device="cuda", dtype=torch.bfloat16)
    flops = 2.0 * T * 6144 * H
    ms = timeit(lambda i: _C.call("b2_gemm_rope", qkv.data_ptr(), xs[i].data_ptr(), w_qkv.data_ptr(), T, 6144, H, H, H,
                                  6144, cos.data_ptr(), sin.data_ptr(), None, 40, T, *nul, st), n_rot)
    r["qkv_rope_fused"] = {"ms": ms, "tflops": flops / ms / 1e9}
    ms = timeit(lambda i: ops.rope_qk(gemm(xs[i], w_qkv, out=qkv), cos, sin, None, 40, 128, T), n_rot)
    r["qkv_then_rope"] = {"ms": ms, "tflops": flops / ms / 1e9}
    for k, v in r.items():
        v["frac_of_measured_peak"] = v["tflops"] / PEAKS["bf16_tflops"]
        print(k, {a: round(b, 3) for a, b in v.items()}, flush=True)


if __name__ == "__main__":
    quick = "--quick" in sys.argv
    results = {"peaks": PEAKS, "gemm": {}, "attention": {}, "elementwise": {}, "fused": {}}
    fns = (bench_gemm, bench_attention, bench_elementwise, bench_fused)
    if "--no-gemm" in sys.argv:
        fns = fns[1:]
    if "--only-gemm" in sys.argv:
        fns = fns[:1]
    if "--only-elementwise" in sys.argv:
        fns = (bench_elementwise,)
    for fn in fns:
        try:
            fn(results, quick) if fn is bench_gemm else fn(results)
        except Exception as e:
            print(f"{fn.__name__} failed: {type(e).__name__}: {e}", flush=True)
            results[fn.__name__ + "_error"] = str(e)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    out = ("kernel_bench_gemm.json" if "--only-gemm" in sys.argv else
           "kernel_bench_elementwise.json" if "--only-elementwise" in sys.argv else "kernel_bench.json")
    json.dump(results, open(os.path.join(ROOT, "gpurun_out", out), "w"), indent=1)
