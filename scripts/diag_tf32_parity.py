"""Print the per-tensor error table of the tf32 whole-model parity cases (tests/test_gpu_tf32_model.py) -- the achieved
figures quoted in DESIGN.md come from this script's output."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from gnnepcsaft_b200 import ops  # noqa: E402
from tests import test_gpu_tf32_model as T  # noqa: E402

torch.set_float32_matmul_precision("highest")
ops.set_matmul_precision("tf32")
for name, cfg, n_graphs, target, _ in T.CASES:
    try:
        res, tc = T.run_tf32_step(cfg, n_graphs, target)
    except AssertionError as e:
        print(f"== {name}: FORCING BAND VIOLATION: {str(e)[:600]}")
        continue
    worst = max(res.items(), key=lambda kv: kv[1][0] / kv[1][1])
    grads = [v[0] for k, v in res.items() if k.startswith("d")]
    print(f"== {name}: depth {T.depth_fwd(cfg)}, tcgen05 launches {tc}, pred {res['pred'][0]:.3e} (bound {res['pred'][1]:.3e}), "
          f"loss {res['loss'][0]:.3e}, grads max {max(grads):.3e} median {sorted(grads)[len(grads) // 2]:.3e} "
          f"(bound {2 * res['pred'][1]:.3e}); worst {worst[0]} at {worst[1][0] / worst[1][1]:.2f} of its bound")
    for k, (e, b) in sorted(res.items(), key=lambda kv: -kv[1][0] / kv[1][1])[:8]:
        print(f"     {k:60s} {e:.3e}  / {b:.3e}")
