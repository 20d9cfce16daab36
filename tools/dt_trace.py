"""Where the single-thread roles of dense_tc_gemm_kernel spend their cycles (needs a FLOWCUDA_BUILD_TRACE=1 build).
Runs one log_prob of a workload eagerly with the trace buffer attached to the LAST gemm launch pattern: the buffer is
overwritten by every GEMM launch, so a workload is cut to ONE flow layer and the kernels are read back after each call
through layer ranges.  usage: python tools/dt_trace.py maf_mnist|maf_made|deep16_nvp"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from normalizing_flows_b200 import _lib, ops
from oracle import flows_oracle as O

wl = sys.argv[1] if len(sys.argv) > 1 else "maf_mnist"
lib = C.CDLL(_lib.LIB_PATH)
buf = torch.zeros(2 * 148 * 8, dtype=torch.int64, device="cuda")  # [full-tile launches | n-half (hidden) launches]
names = ["issuer cycles", "issuer waits accumulator", "issuer waits operands", "entry -> issuer loop", "epilogue g0 waits acc",
         "whole CTA (entry -> exit)", "items", "k stages"]

def report(tag):
    torch.cuda.synchronize()
    for half, what in ((0, tag), (1, tag + " -- LAST n-half (hidden) launch")):
        b = buf.cpu().reshape(2, 148, 8)[half].double()
        live = b[:, 0] > 0
        if not live.any():
            continue
        b = b[live]
        print(f"{what}: CTAs {int(live.sum())}", " | ".join(f"{n} {b[:, i].mean():.0f} (max {b[:, i].max():.0f})" for i, n in enumerate(names)))
    buf.zero_()

if wl == "maf_made":
    lib.flowcuda_debug_set_dt_trace_made(C.c_void_p(buf.data_ptr()))
    sp = O.MadeSpec(784, 1, [512, 512])
    p = O.flatten_layers(O.init_made_params(sp, seed=1234, dtype=torch.float32, scale=0.1)).float().cuda()
    x = torch.randn(16384, 784, device="cuda")
    ms = ops.make_made_spec(784, 1, [512, 512])
    for rep in range(2):
        ops.made_log_prob(ms, p, x)
    report("maf_made 1 layer: LAST gemm = affine Linear")
else:
    lib.flowcuda_debug_set_dt_trace(C.c_void_p(buf.data_ptr()))
    if wl == "maf_mnist":
        so = O.CouplingSpec(784, 1, [512, 512, 512], "rqs", 8, 0.0, 1.0, "uniform"); n = 4096; tr, nb, base = 2, 8, 1
    else:
        so = O.CouplingSpec(784, 1, [512, 512], "affine"); n = 16384; tr, nb, base = 1, 0, 0
    p = O.flatten_layers(O.init_coupling_params(so, seed=1234, dtype=torch.float32, scale=0.1)).float().cuda()
    x = torch.rand(n, 784, device="cuda")
    spec = ops.make_spec(784, 1, so.hidden, tr, nb, 0.0, 1.0, base, _lib.PRECISION_BF16X3)
    for rep in range(2):
        ops.coupling_log_prob(spec, p, x)
    report(f"{wl} 1 layer: LAST gemm = transform Linear")
