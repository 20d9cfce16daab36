"""Print the tcgen05.mma chain timing table (flowcuda_selftest_mma_timing)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C, torch
from normalizing_flows_b200 import _lib
lib = _lib.load_selftest()
out = torch.zeros(4 * 128, dtype=torch.int64, device="cuda")
for rep in range(2):
    _lib.check(lib.flowcuda_selftest_mma_timing(C.c_void_p(out.data_ptr()), out.numel(), None))
    torch.cuda.synchronize()
o = out.cpu().tolist()
print("A_src  N  nacc  n_mma  issue_cyc  total_cyc  cyc/mma")
i = 0
while o[4 * i] != -1:
    N, code, ti, tt = o[4 * i:4 * i + 4]
    a = "smem" if code >= 100000 else "tmem"; code %= 100000
    if code // 1000 in (5, 6, 7):
        a = {7: "bf16 SS, MMAs only", 6: "bf16 SS + fence/commit per stage", 5: "bf16 SS + fence/commit/try_wait per stage"}[code // 1000]
    print(f"{a}  {N:4d}  {code // 1000}  {code % 1000:4d}  {ti:7d}  {tt:7d}  {tt / (code % 1000):7.1f}")
    i += 1
