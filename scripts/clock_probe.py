"""Samples SM clock / power every 20 ms while the student forward runs back to back for ~4 s."""
import os, subprocess, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from geopurify_b200 import synth, ops
from geopurify_b200.student import AffinityPredictor, load_numpy_state
sc = synth.make_scene(n_points=150000, n_views=2, feat_dim=16)
vox = torch.as_tensor(sc.voxel_coords.astype(np.int32)).cuda()
idx = ops.VoxelIndex(vox); nbr27 = idx.kernel_map()
student = load_numpy_state(AffinityPredictor(518, 128, 512), synth.make_student_state())
x = torch.randn(vox.shape[0], student.input_stride, device="cuda"); x[:, 518:] = 0
student.forward_sorted(x, nbr27); torch.cuda.synchronize()
p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap,temperature.gpu",
                      "--format=csv,noheader,nounits", "-lms", "20"], stdout=subprocess.PIPE, text=True)
time.sleep(0.3)
t0 = time.time(); n = 0
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
while time.time() - t0 < 4.0:
    student.forward_sorted(x, nbr27); n += 1
    if n % 8 == 0: torch.cuda.synchronize()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
time.sleep(0.1); p.terminate(); out = p.communicate()[0].strip().splitlines()
rows = [l.split(",") for l in out if l.count(",") >= 3]
clk = np.array([float(r[0]) for r in rows]); pw = np.array([float(r[1]) for r in rows])
cap = sum("Active" in r[2] and "Not" not in r[2] for r in rows)
print(f"student {ms:.3f} ms/forward over {n} runs; samples {len(rows)}; clock MHz min/median/max {clk.min():.0f}/{np.median(clk):.0f}/{clk.max():.0f}; "
      f"power W median/max {np.median(pw):.0f}/{pw.max():.0f}; sw_power_cap active in {cap} samples; temp {rows[-1][3].strip()}")
