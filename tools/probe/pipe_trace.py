"""Debug: timeline of one orbfe_extract_batch call (ORBFE_PIPE_TRACE=1 prints timed events to stderr)."""
import os, sys
os.environ["ORBFE_PIPE_TRACE"] = "1"
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "orb-slam3-noted_b200")]
import torch
import orbfe
B, W, H = 512, 752, 480
ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
frames = torch.from_numpy(np.stack([orbfe.synth.frame(1000 + i % 16, W, H) for i in range(B)])).pin_memory()
cap = ex.cap
kps = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
desc = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
n = np.zeros(B, np.int32); mono = np.zeros(B, np.int32)
import ctypes as C
for it in range(4):
    if it == 3: sys.stderr.write("---- traced call\n")
    st = ex._lib.orbfe_extract_batch(ex._h, C.c_void_p(frames.data_ptr()), B, W, H, W, W * H, 0, 1000, C.c_void_p(kps.data_ptr()),
                                     C.c_void_p(desc.data_ptr()), cap, n.ctypes.data_as(C.c_void_p), mono.ctypes.data_as(C.c_void_p))
    assert st >= 0, st
