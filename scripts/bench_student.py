"""Micro-benchmark of the student forward (tcgen05 path) on one synthetic scene, device-resident.
GEOPURIFY_B200_LIB=<variant .so> selects an A/B build of the library."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from geopurify_b200 import synth, ops
from geopurify_b200.student import AffinityPredictor, load_numpy_state

n = int(sys.argv[1]) if len(sys.argv) > 1 else 150000
sc = synth.make_scene(n_points=n, n_views=2, feat_dim=16)
vox = torch.as_tensor(sc.voxel_coords.astype(np.int32)).cuda()
m = vox.shape[0]
idx = ops.VoxelIndex(vox)
nbr27 = idx.kernel_map()
pairs = int((nbr27 >= 0).sum())
student = load_numpy_state(AffinityPredictor(518, 128, 512), synth.make_student_state())
x = torch.randn(m, student.input_stride, device="cuda")
x[:, 518:] = 0
ref = None
for _ in range(2):
    e = student.forward_sorted(x, nbr27)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
R = 5
for _ in range(R):
    e = student.forward_sorted(x, nbr27)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / R
fl = 2.0 * (pairs * 544 * 512 + 8 * pairs * 512 * 512 + m * 512 * 128)
print(f"lib={os.environ.get('GEOPURIFY_B200_LIB','default')} M={m} pairs={pairs} student {ms:.3f} ms  {fl/ms/1e9:.1f} TFLOP/s useful  checksum {e.double().sum().item():.6f}")
