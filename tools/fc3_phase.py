"""globaltimer stamps (ns) of CTA 0 of the fused fc3 cluster kernel, run inside the engine's training step."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.ops import native, functional as K
L = native.lib()
L.ddl_debug_stamps.argtypes = [C.c_int, C.c_void_p, C.c_int]
dev = torch.device("cuda")
B, Kin, NC = 100, 512, 10
h = torch.randn(B, Kin, device=dev); w = torch.randn(Kin, NC, device=dev) * .1; b = torch.randn(NC, device=dev)
y = torch.randint(0, NC, (B,), device=dev, dtype=torch.int32)
st = native.new_step_state(seed=1)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
L.ddl_debug_stamps(1, None, 0)
names = ["start", "phase0 issued (w->smem, bias, labels)", "pdl_wait done", "syncthreads (swT complete)", "cluster wait 1",
         "logits+softmax+push done", "shs stored", "cluster barrier 2", "dW done", "db done", "dh done", "end"]
for cold in (True, False):
    for _ in range(3):
        if cold:
            flush.fill_(1)
        a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); K.fc3_fused(h, w, b, y, st, keep_prob=0.5); e.record(); torch.cuda.synchronize()
    buf = (C.c_uint64 * 16)()
    L.ddl_debug_stamps(1, buf, 16)
    print("[%s] " % ("cold" if cold else "warm") + " | ".join("%s +%.2f" % (names[i], (buf[i] - buf[0]) / 1e3) for i in range(1, 12)))
