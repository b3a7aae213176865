"""Which combination breaks cudaMalloc with 'resource already mapped'?  python scripts/debug_sharded_pinned.py <case>"""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ompl_b200 as ob
from ompl_b200._lib import check, lib
case = sys.argv[1]
L = lib()
rt = torch.cuda.cudart()
pts, qs = np.random.default_rng(1).random((200000, 7)), np.random.default_rng(2).random((4096, 7))
ids = np.empty((4096, 10), dtype=np.uint64); d = np.empty((4096, 10)); cnt = np.empty(4096, dtype=np.uint32)
def pin(a, flags=0):
    rc = rt.cudaHostRegister(a.ctypes.data, a.nbytes, flags); print("  register rc", rc, flush=True)
if case in ("torch_first", "torch_first_pin"):
    torch.zeros(1, device="cuda:0"); torch.cuda.synchronize()
if case == "pin_before":
    torch.zeros(1, device="cuda:0")
    for a in (qs, ids, d): pin(a)
nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(7), devices=[0, 1])
nn.add_rows(pts)
if case in ("pin_after", "torch_first_pin"):
    for a in (qs, ids, d): pin(a)
if case == "pin_portable":
    for a in (qs, ids, d): pin(a, 1)
for it in range(3):
    check(L.omplcu_nn_knn(nn.handle, qs.ctypes.data, 4096, 10, ids.ctypes.data, d.ctypes.data, cnt.ctypes.data))
print(case, "ok", ids[0, :3])
