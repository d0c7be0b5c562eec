import torch, sys
sys.path.insert(0, '.')
from adkf_ift_b200 import ops
dev = torch.device('cuda:0')
M, K, N = 700, 512, 2048
wfull = torch.randn(N, K + 2048, device=dev)
dy = torch.randn(M, N, device=dev)
x = torch.randn(M, K, device=dev)
for name, fn in [("strided KT", lambda: ops.gemm_tf32x3(dy, False, wfull[:, :K], True)),
                 ("contig KT", lambda: ops.gemm_tf32x3(dy, False, wfull[:, :K].contiguous(), True)),
                 ("TT", lambda: ops.gemm_tf32x3(dy, True, x, True)),
                 ("strided fwd", lambda: ops.gemm_tf32x3(x, False, wfull[:, :K], False))]:
    try:
        o = fn(); torch.cuda.synchronize(); print(name, "ok", o.shape)
    except Exception as e:
        print(name, "FAIL", e)
ref = dy.double() @ wfull[:, :K].double()
try:
    o = ops.gemm_tf32x3(dy, False, wfull[:, :K], True)
    print("err", float((o.double() - ref).abs().max() / ref.abs().max()))
except Exception as e:
    print("x", e)
