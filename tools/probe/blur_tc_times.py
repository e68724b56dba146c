"""Debug (needs a library built with -DBT_TIMING=1): where the roles of k_blur_tc wait, in cycles per tile (CTA 0, lane 0 of each warp).
   slots: 0 producer<-pEmpty (x4 warps)  1 MMA<-pFull  2 MMA<-d1Empty  3 MMA<-hFull  4 MMA<-d2Empty
          5 epi1<-d1Full (x4)  6 epi1<-hEmpty (x4)  7 epi2<-d2Full (x4)  8 total  9 tiles"""
import ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "orb-slam3-noted_b200")]
import torch
import orbfe
from orbfe import _lib
lib = _lib.lib()
B, W, H = 128, 752, 480
ex = orbfe.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
ex.set_blur_backend(ex.BLUR_TENSOR)
frames = torch.from_numpy(np.stack([orbfe.synth.frame(1000 + i % 16, W, H) for i in range(B)])).cuda()
for _ in range(2):
    ex.extract_batch_device(frames.data_ptr(), B, W, H)
ex.synchronize()
t = (C.c_longlong * 16)()
lib.orbfe_debug_blur_tc_times(t, 1)
ex.extract_batch_device(frames.data_ptr(), B, W, H)
ex.synchronize()
lib.orbfe_debug_blur_tc_times(t, 0)
v = list(t)
n = max(v[9], 1)
names = ["prod<-pEmpty/4", "mma<-pFull", "mma<-d1Empty", "mma<-hFull", "mma<-d2Empty", "epi1<-d1Full/4", "epi1<-hEmpty/4", "epi2<-d2Full/4"]
div = [4, 1, 1, 1, 1, 4, 4, 4]
print("tiles of CTA 0:", v[9], " cycles per tile:", v[8] / n)
for i, nm in enumerate(names):
    print(f"  {nm:16s} {v[i] / div[i] / n:8.0f} cycles per tile")
