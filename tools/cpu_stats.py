import sys; sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
import numpy as np, ctypes as C, sys
from mpi_cuda_largescaleknn_b200.ops._lib import lib
from mpi_cuda_largescaleknn_b200.utils.oracle import cpu_knn, bruteforce, bit_equal
L=lib()
n=int(sys.argv[1]) if len(sys.argv)>1 else 400000
x=np.random.default_rng(0).random((n,3),dtype=np.float32)
st=(C.c_uint64*4)()
L.b200knn_cpu_stats(st,1)
for k in (100,10):
    r=cpu_knn(x,k=k)
    L.b200knn_cpu_stats(st,1)
    g=(len(x)+31)//32
    print("k",k,"leaves/group %.1f inserts/query %.1f inner/group %.1f"%(st[0]/g,st[2]/len(x),st[3]/g))
sel=np.arange(0,n,n//500)
print(bit_equal(cpu_knn(x,k=100)[sel], bruteforce(x,queries=x[sel],k=100)))
