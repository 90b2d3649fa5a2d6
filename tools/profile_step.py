"""One factorize + inertia + solve of a bench workload between cudaProfilerStart/Stop, for ncu launch lists:

  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv \
      python tools/profile_step.py <workload> [option=value ...]

Without ncu it prints the CUDA-event times of the same step (graphs on), which is what the shares are compared with."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

import bench
from uno_b200 import KKTSolver, generators as G

name = sys.argv[1]
kkt = bench.WORKLOADS[name][1](G, 0)
s = KKTSolver(device=0)
for kv in sys.argv[2:]:
    k, v = kv.split("=")
    s.set_option(k, float(v))
s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)
d_val = torch.from_numpy(kkt.val).cuda()
d_rhs = torch.from_numpy(kkt.rhs).cuda()
d_x = torch.empty_like(d_rhs)
for _ in range(3):
    s.factorize_device(d_val.data_ptr())
    s.solve_device(d_rhs.data_ptr(), d_x.data_ptr())
torch.cuda.synchronize()
l0 = s.stats()["kernel_launches"]
torch.cuda.profiler.start()
s.factorize_device(d_val.data_ptr())
berr = s.solve_device(d_rhs.data_ptr(), d_x.data_ptr())
torch.cuda.synchronize()
torch.cuda.profiler.stop()
st = s.stats()
print(json.dumps({"workload": name, "N": kkt.dim, "factor_ms": 1e3 * st["t_factor_last"], "solve_ms": 1e3 * st["t_solve_last"], "refine_steps": st["refine_steps_last"],
                  "launches": st["kernel_launches"] - l0, "inertia": s.get_inertia(), "backward_error": berr, "componentwise": st["backward_error_componentwise_last"],
                  "fronts": st["num_fronts"], "levels": st["num_levels"], "factor_flops": st["flops_factor"], "nnz_factor": st["nnz_factor"]}))
