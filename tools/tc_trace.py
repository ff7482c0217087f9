import os; os.environ.setdefault("DDL_LAB", "1")   # needs the instrumented library (ops/build.py --lab)
import os, sys, ctypes as C
os.environ["DDL_TC_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.ops import native
L = native.lib()
L.ddl_tc_plan_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
dev = torch.device("cuda")
N, cred, nout = 128, 2048, 32
x = torch.randn(N, cred, device=dev); w = torch.randn(cred * nout, device=dev); out = torch.zeros(N, nout, device=dev)
plan = L.ddl_tc_plan_act(0, x.data_ptr(), w.data_ptr(), out.data_ptr(), N, 1, 1, cred, nout, 1, 32, 1, 0, None, None, 1.0, None, 0, 1.0, None, None)
for _ in range(3):
    L.ddl_tc_plan_run(plan, native.stream_ptr())
torch.cuda.synchronize()
buf = (C.c_uint64 * (64 * 4))()
L.ddl_tc_plan_trace(plan, buf, 64 * 4)
t0 = buf[0]
print("iter: prod_after_empty_wait | mma_before_wait  mma_after_wait  mma_after_commit  (cycles rel.)")
for i in range(40):
    print(i, buf[i*4]-t0, "|", buf[i*4+1]-t0, buf[i*4+2]-t0, buf[i*4+3]-t0, " wait=%d issue=%d" % (buf[i*4+2]-buf[i*4+1], buf[i*4+3]-buf[i*4+2]))
