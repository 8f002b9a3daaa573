"""Where does the single-LP latency go?  Host enqueue time vs device time for B = 1 tiny problems."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from linear_programming_b200 import _lib, workloads
L = _lib.lib()
for (n, m) in ((3, 3), (50, 100), (100, 100)):
    A, b, c = workloads.simplex_batch("D1", 1, n, m, seed=2)
    dA, db, dc = (torch.from_numpy(v).cuda() for v in (A, b, c))
    st = torch.empty(1, dtype=torch.int32, device="cuda"); x = torch.empty((1, m), dtype=torch.float64, device="cuda")
    ob = torch.empty(1, dtype=torch.float64, device="cuda"); npv = torch.empty(1, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    def call():
        _lib.check(L.lpb200_simplex_batch_device(1, n, m, dA.data_ptr(), db.data_ptr(), dc.data_ptr(), 1e-9, 100000,
                                                 st.data_ptr(), x.data_ptr(), ob.data_ptr(), npv.data_ptr(), None, 0, stream))
    for _ in range(20): call()
    torch.cuda.synchronize()
    R = 200
    t0 = time.perf_counter()
    for _ in range(R): call()
    t_enq = (time.perf_counter() - t0) / R
    torch.cuda.synchronize()
    t_all = (time.perf_counter() - t0) / R
    # one call, device time
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record(); call(); e1.record(); torch.cuda.synchronize()
    print("n=%d m=%d pivots=%d  enqueue %.1f us/call  back-to-back %.1f us/call  single call device %.1f us"
          % (n, m, int(npv.item()), t_enq * 1e6, t_all * 1e6, e0.elapsed_time(e1) * 1e3), flush=True)
