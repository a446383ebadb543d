#!/usr/bin/env python
"""Per-kernel roofline table: every hot-path kernel called alone through the C ABI at a
fine-level size, timed with CUDA events on the launching stream (inputs >> L2), reported as
algorithmic GB/s (SURVEY.md 8(d) bytes per fine point) over the measured HBM copy peak.

  python tools/fine_kernels.py [--n2 4097] [--n3 385] [--reps 10] [--only 2d|3d]
With --reps 1 --warm 0 it is the target of the `ncu --set full` captures in profiles/."""
import argparse
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cedar_b200.abi as abi

L = abi.lib
PK = None


def P(t):
    return C.c_void_p(t.data_ptr())


def peak():
    try:
        return float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def emit(out, pk, **o):
    o["gbs"] = o["bytes_per_pt"] * o["pts"] / max(o["ms"], 1e-6) / 1e6
    o["frac_measured_peak"] = o["gbs"] / pk
    out.append(o)
    print(json.dumps(o), flush=True)


def timeit(fn, reps, warm):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = C.c_void_p(), C.c_void_p()
    L.bmgx_event_create(C.byref(e0)); L.bmgx_event_create(C.byref(e1))
    l0 = L.bmgx_launch_count()
    L.bmgx_event_record(e0, None)
    for i in range(reps):
        fn(i)
    L.bmgx_event_record(e1, None)
    L.bmgx_event_sync(e1)
    nl = (L.bmgx_launch_count() - l0) / reps
    ms = C.c_float()
    L.bmgx_event_elapsed_ms(e0, e1, C.byref(ms))
    L.bmgx_event_destroy(e0); L.bmgx_event_destroy(e1)
    return ms.value / reps, nl


PK = peak()


def rnd(*shape):
    return torch.rand(shape, dtype=torch.float64, device="cuda") + 0.5


def coarse_n(n):
    return (n - 1) // 2 + 1


def table2(n, reps, warm, out):
    II = n + 2
    nc = coarse_n(n); IIC = nc + 2
    pts = n * n
    for nine in (True, False):
        nst = 5 if nine else 3
        ifd = 0 if nine else 1
        tag = "9pt" if nine else "5pt"
        so = rnd(nst, II, II); so[0] += 12.0
        for t in (so,):
            t[:, 0, :] = 0; t[:, -1, :] = 0; t[:, :, 0] = 0; t[:, :, -1] = 0
        q = rnd(II, II); q2 = rnd(II, II); b = rnd(II, II); r = torch.zeros(II, II, dtype=torch.float64, device="cuda")
        sor = torch.zeros(2, II, II, dtype=torch.float64, device="cuda")
        ci = rnd(8, IIC, IIC) * 0.25
        qc = rnd(IIC, IIC)
        soc = torch.zeros(5, IIC, IIC, dtype=torch.float64, device="cuda")
        abi.check(L.bmgx2_SETUP_recip(P(so), P(sor), II, II, nst, 2, None))
        bufs = [q, q2]
        K = []
        K.append((f"relax_GS_oop {tag} (fused sweep)", 64 if nine else 48,
                  lambda i: L.bmgx2_relax_GS_oop(1, P(so), P(b), P(bufs[i & 1]), P(bufs[(i + 1) & 1]), P(sor), II, II, 1, ifd, nst, 2, 1, i & 1, 0, None)))
        K.append((f"relax_GS {tag} (in place, per colour)", 64 if nine else 48,
                  lambda i: L.bmgx2_relax_GS(1, P(so), P(b), P(q), P(sor), II, II, 1, ifd, nst, 2, 1, i & 1, 0, None)))
        K.append((f"residual {tag}", 64 if nine else 48,
                  lambda i: L.bmgx2_residual(0, P(so), P(b), P(q), P(r), II, II, 0, ifd, nst, 0, 0, 0, 0, None)))
        if nine:
            K.append(("restrict", 26, lambda i: L.bmgx2_restrict(P(r), P(qc), P(ci), II, II, IIC, IIC, 0, None)))
        K.append((f"interp_add {tag}", 50,
                  lambda i: L.bmgx2_interp_add(P(q), P(qc), P(r), P(so), P(ci), IIC, IIC, II, II, nst, 0, None)))
        K.append((f"SETUP_interp_OI {tag}", 56 if nine else 40,
                  lambda i: L.bmgx2_SETUP_interp_OI(P(so), P(soc), P(ci), II, II, IIC, IIC, ifd, nst, 0, 0, None)))
        K.append((f"SETUP_ITLI_ex {tag}", 66 if nine else 50,
                  lambda i: L.bmgx2_SETUP_ITLI_ex(P(so), P(soc), P(ci), II, II, IIC, IIC, ifd, nst, 0, None)))
        K.append((f"SETUP_recip {tag}", 16, lambda i: L.bmgx2_SETUP_recip(P(so), P(sor), II, II, nst, 2, None)))
        K.append((f"sumsq (l2 norm)", 8, lambda i: L.bmgx2_sumsq(P(q), II, II, P(wk), P(o1), None)))
        wk = torch.zeros(L.bmgx_norm_work_len(), dtype=torch.float64, device="cuda")
        o1 = torch.zeros(1, dtype=torch.float64, device="cuda")
        for name, bpp, fn in K:
            ms, nl = timeit(lambda i: abi.check(fn(i)), reps, warm)
            emit(out, PK, kernel="2D " + name, n=n, pts=pts, bytes_per_pt=bpp, ms=ms, launches=nl)
        # lines
        sorl = torch.zeros(2, II, II, dtype=torch.float64, device="cuda")
        scratch = torch.zeros(II, II, dtype=torch.float64, device="cuda")
        for d, setup, relax in (("x", L.bmgx2_SETUP_lines_x, L.bmgx2_relax_lines_x), ("y", L.bmgx2_SETUP_lines_y, L.bmgx2_relax_lines_y)):
            ms, nl = timeit(lambda i: abi.check(setup(P(so), P(sorl), II, II, nst, 0, None)), max(1, reps // 2), min(warm, 1))
            emit(out, PK, kernel=f"2D SETUP_lines_{d} {tag}", n=n, pts=pts, bytes_per_pt=32, ms=ms, launches=nl)
            ms, nl = timeit(lambda i: abi.check(relax(1, P(so), P(b), P(q), P(sorl), P(scratch), II, II, 1, ifd, nst, 1, i & 1, 0, None)), reps, warm)
            bpp = 64 if nine else 48
            emit(out, PK, kernel=f"2D relax_lines_{d} {tag}", n=n, pts=pts, bytes_per_pt=bpp, ms=ms, launches=nl)
        del so, q, q2, b, r, sor, ci, qc, soc, sorl, scratch, bufs, K
        torch.cuda.empty_cache()


def table3(n, reps, warm, out):
    II = n + 2
    nc = coarse_n(n); IIC = nc + 2
    pts = n ** 3
    for full in (False, True):
        nst = 14 if full else 4
        ifd = 0 if full else 1
        tag = "27pt" if full else "7pt"
        so = rnd(nst, II, II, II); so[0] += 40.0
        so[:, 0] = 0; so[:, -1] = 0; so[:, :, 0] = 0; so[:, :, -1] = 0; so[:, :, :, 0] = 0; so[:, :, :, -1] = 0
        q = rnd(II, II, II); b = rnd(II, II, II); r = torch.zeros_like(q)
        sor = torch.zeros(2, II, II, II, dtype=torch.float64, device="cuda")
        ci = rnd(26, IIC, IIC, IIC) * 0.1
        qc = rnd(IIC, IIC, IIC)
        soc = torch.zeros(14, IIC, IIC, IIC, dtype=torch.float64, device="cuda")
        abi.check(L.bmgx3_SETUP_recip(P(so), P(sor), II, II, II, nst, 2, None))
        K = []
        K.append((f"relax_GS {tag}", 136 if full else 56,
                  lambda i: L.bmgx3_relax_GS(1, P(so), P(b), P(q), P(sor), II, II, II, ifd, nst, 2, 1, i & 1, 0, None)))
        if not full:
            q2 = rnd(II, II, II)
            bufs = [q, q2]
            K.append((f"relax_GS_oop {tag} (fused sweep)", 56,
                      lambda i: L.bmgx3_relax_GS_oop(1, P(so), P(b), P(bufs[i & 1]), P(bufs[(i + 1) & 1]), P(sor), II, II, II, ifd, nst, 2, 1, i & 1, 0, None)))
        K.append((f"residual {tag}", 136 if full else 56,
                  lambda i: L.bmgx3_residual(1, 1, ifd, P(q), P(b), P(so), P(r), II, II, II, nst, None)))
        if full:
            K.append(("restrict", 35, lambda i: L.bmgx3_restrict(P(r), P(qc), P(ci), II, II, II, IIC, IIC, IIC, 0, None)))
        K.append((f"interp_add {tag}", 59,
                  lambda i: L.bmgx3_interp_add(P(q), P(qc), P(so), P(r), P(ci), IIC, IIC, IIC, II, II, II, nst, 0, None)))
        K.append((f"SETUP_interp_OI {tag}", 138 if full else 58,
                  lambda i: L.bmgx3_SETUP_interp_OI(P(so), P(soc), P(ci), II, II, II, IIC, IIC, IIC, ifd, nst, 0, 0, None, None)))
        itli = L.bmgx3_SETUP_ITLI27_ex if full else L.bmgx3_SETUP_ITLI07_ex
        K.append((f"SETUP_ITLI{'27' if full else '07'}_ex", 152 if full else 72,
                  lambda i: itli(P(so), P(soc), P(ci), II, II, II, IIC, IIC, IIC, 0, None)))
        for name, bpp, fn in K:
            ms, nl = timeit(lambda i: abi.check(fn(i)), reps, warm)
            emit(out, PK, kernel="3D " + name, n=n, pts=pts, bytes_per_pt=bpp, ms=ms, launches=nl)
        del so, q, b, r, sor, ci, qc, soc, K
        torch.cuda.empty_cache()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--n2", type=int, default=4097)
    ap.add_argument("--n3", type=int, default=385)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--warm", type=int, default=2)
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    pk = PK
    out = []
    if a.only != "3d":
        table2(a.n2, a.reps, a.warm, out)
    if a.only != "2d":
        table3(a.n3, a.reps, a.warm, out)
    print(f"# peak {pk} GB/s", flush=True)
    print("| kernel | n | B/pt | ms | launches | GB/s | % measured peak |\n|---|---|---|---|---|---|---|")
    for o in out:
        print(f"| {o['kernel']} | {o['n']} | {o['bytes_per_pt']} | {o['ms']:.3f} | {o['launches']:.0f} | {o['gbs']:.0f} | {100 * o['frac_measured_peak']:.0f} |")
