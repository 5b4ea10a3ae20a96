"""Golden track tables from the REFERENCE's own code, run in the build container (where /root/reference is mounted):
python/motion_planning.py is imported unmodified, with stand-ins for the packages it imports but which are absent here
(matplotlib, casadi, track_database: not used by the functions called; qpsolvers.solve_qp: the spline-fit QP has
equality constraints only, so its exact solution is one KKT solve, done here with scipy -- the reference passes it to
proxqp).  Writes tests/golden/tracks_reference.npz: for every packaged track the tables of offline_motion_plan
(motion_planning.py:345-399).

  python tests/golden/make_track_golden.py [/root/reference]
"""
import os
import sys
import types

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

HERE = os.path.dirname(os.path.abspath(__file__))
TRACKS = ("fsds_default", "fsds_competition_1", "fsds_competition_2", "fsds_competition_3", "skidpad", "short_skidpad")


def import_reference_motion_planning(ref_root: str):
    """The reference module with its missing third-party imports replaced by inert stand-ins."""
    def solve_qp(P, q, A=None, b=None, solver=None, **kw):     # min 1/2 p'Pp + q'p  s.t.  A p = b
        P, A = sp.csc_matrix(P), sp.csc_matrix(A)
        n, m = P.shape[0], A.shape[0]
        K = sp.bmat([[P, A.T], [A, None]], format="csc")
        return spla.spsolve(K, np.concatenate([-np.asarray(q).ravel(), np.asarray(b).ravel()]))[:n]

    qps = types.ModuleType("qpsolvers"); qps.solve_qp = solve_qp; qps.available_solvers = ["proxqp"]
    mpl = types.ModuleType("matplotlib"); plt = types.ModuleType("matplotlib.pyplot"); mpl.pyplot = plt
    cas = types.ModuleType("casadi")
    for n_ in ("MX", "SX", "exp", "tanh"):
        setattr(cas, n_, object)
    tdb = types.ModuleType("track_database"); tdbu = types.ModuleType("track_database.utils")

    class Track:   # what motion_planning.py reads of track_database.Track: center_line [n,2], track_widths [n,2]
        def __init__(self, name):
            rows = []
            with open(os.path.join(ref_root, "data", name, "center_line.csv")) as f:
                for i, line in enumerate(f):
                    line = line.strip()
                    if not line or (i == 0 and (line.startswith("#") or line[0].isalpha())):
                        continue
                    rows.append([float(v) for v in line.split(",")[:4]])
            arr = np.asarray(rows)
            self.center_line, self.track_widths = arr[:, :2].copy(), arr[:, 2:4].copy()

    tdb.Track = Track; tdbu.plot_cones = lambda *a, **k: None; tdb.utils = tdbu
    for name, mod in (("qpsolvers", qps), ("matplotlib", mpl), ("matplotlib.pyplot", plt), ("casadi", cas),
                      ("track_database", tdb), ("track_database.utils", tdbu)):
        sys.modules.setdefault(name, mod)
    sys.path.insert(0, os.path.join(ref_root, "python"))
    import motion_planning   # noqa: E402  (the reference's module)
    return motion_planning


def reference_tables(ref_root: str):
    mpmod = import_reference_motion_planning(ref_root)
    argv, sys.argv = sys.argv, sys.argv[:1]        # offline_motion_plan reads the QP solver's name from sys.argv[1]
    out = {}
    try:
        for name in TRACKS:
            mp = mpmod.offline_motion_plan(name)
            for f in ("s_ref", "X_ref", "Y_ref", "phi_ref", "kappa_ref", "right_widths", "left_widths"):
                out[f"{name}/{f}"] = np.asarray(getattr(mp, f), dtype=np.float64)
            out[f"{name}/lap_length"] = np.float64(mp.lap_length)
            t3 = mpmod.triple_motion_plan_ref(mp)
            out[f"{name}/s_ref_tripled"] = np.asarray(t3.s_ref, dtype=np.float64)
    finally:
        sys.argv = argv
    return out


if __name__ == "__main__":
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    tabs = reference_tables(ref)
    np.savez_compressed(os.path.join(HERE, "tracks_reference.npz"), **tabs)
    # the CSV the reference writes for a track (motion_planning.py:431-456), one track as a fixture
    mpmod = import_reference_motion_planning(ref)
    argv, sys.argv = sys.argv, sys.argv[:1]
    mpmod.generate_track_data_file("fsds_default", os.path.join(HERE, "fsds_default_track_data_reference.csv"))
    sys.argv = argv
    for name in TRACKS:
        print(f"{name:22s} lap_length {float(tabs[name + '/lap_length']):10.4f}  kappa in [{tabs[name + '/kappa_ref'].min():+.4f}, {tabs[name + '/kappa_ref'].max():+.4f}]")
