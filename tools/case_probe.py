"""Debug helper: convert ONE frame on the GPU and compare with the oracle. usage: case_probe.py fmt w h seed"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import formulas as F
from video_edge_ai_proxy_b200 import FramePath, desc_from_tensors
from video_edge_ai_proxy_b200.synth import make_planes
fmt, w, h, seed = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
planes = make_planes(fmt, w, h, seed)
path = FramePath(0)
t = [torch.from_numpy(p).cuda() for p in planes]
out = torch.full((h, w, 3), 0xA5, dtype=torch.uint8, device="cuda")
path.convert_batch([desc_from_tensors(fmt, t, out)], torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
ok = np.array_equal(out.cpu().numpy(), F.to_bgr24(fmt, planes))
print(fmt, w, h, "OK" if ok else "MISMATCH")
