"""Time one NT GEMM of the DMMA kernel for the raster group the environment names (GPK_GEMM_RASTER=G tile-rows per
group, default 8; read once per process, so run this once per G).  Shapes: the 8192^3 launch `roofline.traffic` is
quoted on, and the two largest launches of an N=20k evaluation (LAUUM 20096^2 x 20096 both-upper lower-only; the
root trailing update 10112^2 x 9984 lower-only).  Timing: CUDA events, best of 5 after 2 warm-ups.  For the DRAM
bytes run the same command under `ncu --set full --clock-control none -k regex:gpk_gemm_nt -c 3` and read
dram__bytes_read.sum + dram__bytes_write.sum per launch (tools/README.md); only the SHARE of re-reads is compared
there, never the time."""
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200 import _lib  # noqa: E402

lib = _lib.load()
P = lambda t: C.c_void_p(t.data_ptr())
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
G = os.environ.get("GPK_GEMM_RASTER", "8")


def run(name, M, N, K, args, flops):
    a = torch.randn(M, K, dtype=torch.float64, device="cuda")
    b = a if N == M else torch.randn(N, K, dtype=torch.float64, device="cuda")
    c = torch.zeros(M, N, dtype=torch.float64, device="cuda")
    best = 1e30
    for i in range(7):
        e0.record()
        rc = lib.gpk_gemm_nt(P(a), K, P(b), K, P(c), N, M, N, K, *args, st)
        e1.record()
        torch.cuda.synchronize()
        assert rc == 0, rc
        if i >= 2:
            best = min(best, e0.elapsed_time(e1))
    print(json.dumps({"raster_rows": int(G), "shape": name, "M": M, "N": N, "K": K, "ms": round(best, 3),
                      "tflops_algorithmic": round(flops / best / 1e9, 2)}), flush=True)
    del a, b, c


# (alpha, accumulate, a_tri, b_tri, lower_only): 0 none, 1 lower, 2 upper
run("plain_8192", 8192, 8192, 8192, (1.0, 0, 0, 0, 0), 2.0 * 8192 ** 3)
if "--small" not in sys.argv:
    n = 20096
    run("lauum_20096", n, n, n, (1.0, 0, 2, 2, 1), n ** 3 / 3.0)
    run("syrk_root_10112x9984", 10112, 10112, 9984, (-1.0, 1, 0, 0, 1), 10112.0 ** 2 * 9984)
