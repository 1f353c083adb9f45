import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga
m, n, k = int(sys.argv[1]), 4096, 7168
out_dtype = torch.float32 if sys.argv[2] == "f32" else torch.bfloat16
fields = {kv.split("=")[0]: int(kv.split("=")[1]) for kv in sys.argv[3:]}
a = torch.randn((m, k), device="cuda", dtype=torch.bfloat16)
w = torch.randn((n, k), device="cuda", dtype=torch.bfloat16)
c = torch.empty((m, n), device="cuda", dtype=out_dtype)
dga.matmul_with_plan(a, w.t(), c, **fields)
torch.cuda.synchronize()
ref = a.float() @ w.float().t()
print(sys.argv[1:], "ok maxerr", float((c.float() - ref).abs().max() / ref.abs().max()))
