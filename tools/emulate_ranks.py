"""per-rank work of an N-GPU run, measured on ONE GPU: builds rank r's share of bench.py's problem (owner partition) for
r = 0..N-1 in turn and times its kernels.  usage: python tools/emulate_ranks.py [N] [nx]"""
import ctypes, json, os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import __graft_entry__ as ge
ge.build()
from ngsxfem_b200 import _lib
L = _lib.load()
N = int(sys.argv[1]) if len(sys.argv) > 1 else 8
nx = int(sys.argv[2]) if len(sys.argv) > 2 else 119
rows = []
for r in range(N):
    P = bench.build_problem(nx, "cuda:0", r, N, 6)
    mesh, lsetp1, ci, a, xfem, mesh_int = P["mesh"], P["lsetp1"], P["ci"], P["a"], P["xfem"], P["mesh_int"]
    a.Assemble()
    torch.cuda.synchronize()
    plan = a._plans[0]
    ncut = int((ci.dom[plan.els.long()] == 2).sum()) if ci.dom.numel() == mesh.ne else -1
    ev = lambda: torch.cuda.Event(enable_timing=True)
    L.ngsx_debug_set_timing(1)
    tb = (ctypes.c_double * 3)()
    st = {"classify": [], "assemble": [], "integrate": []}
    for it in range(8):
        e = [ev() for _ in range(4)]
        e[0].record(); ci.Update(lsetp1); e[1].record(); a.Assemble(); e[2].record()
        xfem.IntegrateX(lsetp1, mesh_int, cf=1.0, order=4, domain_type=xfem.IF, sync=False)
        xfem.IntegrateX(lsetp1, mesh_int, cf=1.0, order=4, domain_type=xfem.NEG, sync=False); e[3].record()
        torch.cuda.synchronize()
        if it >= 3:
            st["classify"].append(e[0].elapsed_time(e[1])); st["assemble"].append(e[1].elapsed_time(e[2])); st["integrate"].append(e[2].elapsed_time(e[3]))
    L.ngsx_debug_get_timing(tb)
    L.ngsx_debug_set_timing(0)
    full = xfem.CutInfo(mesh, lsetp1)
    ncut = int((full.dom[plan.els.long()] == 2).sum())
    rows.append({"rank": r, "listed": plan.els.numel(), "cut_listed": ncut, "pairs": getattr(plan, "npairs", None),
                 "k_geom": tb[0], "k_cutmat": tb[1], "k_rows": tb[2], **{k: statistics.mean(v) for k, v in st.items()}})
    print(json.dumps(rows[-1]))
    del P, a, ci, plan, full
    torch.cuda.empty_cache()
