"""Generate ihm2_b200/data/tracks/<name>.npz (single-lap motion-plan tables + the raw centre
line) from the reference's data/<name>/center_line.csv.  Run in the build container, where
/root/reference is mounted; the GPU box only sees the generated files."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200.track import PACKAGED_TRACKS, load_center_line, offline_motion_plan  # noqa: E402

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/data"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "ihm2_b200", "data", "tracks")
os.makedirs(OUT, exist_ok=True)
for name in PACKAGED_TRACKS:
    cl, w = load_center_line(os.path.join(REF, name, "center_line.csv"))
    mp = offline_motion_plan(cl, w)
    np.savez_compressed(
        os.path.join(OUT, f"{name}.npz"),
        s_ref=mp.s_ref, X_ref=mp.X_ref, Y_ref=mp.Y_ref, phi_ref=mp.phi_ref, kappa_ref=mp.kappa_ref,
        right_widths=mp.right_widths, left_widths=mp.left_widths, lap_length=mp.lap_length,
        center_line=cl, track_widths=w,
    )
    print(f"{name:22s} n={len(cl):4d} lap_length={mp.lap_length:9.4f} "
          f"kappa in [{mp.kappa_ref.min():+.4f}, {mp.kappa_ref.max():+.4f}]")
