"""host-side cost of one bench step on a tiny mesh (launch-bound): wall time per call + cProfile over 300 steps"""
import cProfile, pstats, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import __graft_entry__ as ge
ge.build()
P = bench.build_problem(int(sys.argv[1]) if len(sys.argv) > 1 else 16, "cuda:0")
mesh, lsetp1, ci, a, xfem, mesh_int = P["mesh"], P["lsetp1"], P["ci"], P["a"], P["xfem"], P["mesh_int"]
a.Assemble()
torch.cuda.synchronize()

def step():
    ci.Update(lsetp1)
    a.Assemble()
    s_if = xfem.IntegrateX(lsetp1, mesh_int, cf=1.0, order=4, domain_type=xfem.IF, sync=False)
    s_neg = xfem.IntegrateX(lsetp1, mesh_int, cf=1.0, order=4, domain_type=xfem.NEG, sync=False)
    return s_if, s_neg

for _ in range(20):
    step()
torch.cuda.synchronize()
for name, fn in (("Update", lambda: ci.Update(lsetp1)), ("Assemble", a.Assemble),
                 ("Integrate", lambda: xfem.IntegrateX(lsetp1, mesh_int, cf=1.0, order=4, domain_type=xfem.IF, sync=False)), ("step", step)):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(300):
        fn()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"{name}: host issue {1e6 * (t1 - t0) / 300:.1f} us/call, with drain {1e6 * (t2 - t0) / 300:.1f} us/call")
pr = cProfile.Profile()
pr.enable()
for _ in range(300):
    step()
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
