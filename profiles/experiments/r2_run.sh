set -x
timeout 600 python -m pytest tests/test_gpu_dp.py -q -m gpu -x -k emulated 2>&1 | grep -E "^>|^E  " | head -12
python - <<'PY'
import sys; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import torch, hcatgnet_b200 as pkg
import tests.test_gpu_dp as T
whole, sh_a, sh_b = T._shards(pkg)
sd0 = {k: v.clone() for k, v in T._fresh(pkg).state_dict().items()}
one = pkg.FusedTrainer(T._fresh(pkg, sd0)); ra, rb = pkg.FusedTrainer(T._fresh(pkg, sd0)), pkg.FusedTrainer(T._fresh(pkg, sd0))
for it in range(4):
    l1 = one.step(whole.to("cuda")); ra.accumulate(sh_a.to("cuda")); rb.accumulate(sh_b.to("cuda"))
    total = ra.grad + rb.grad
    gd = (total[:-2]-one.grad[:-2]).abs().max().item(); gm = one.grad[:-2].abs().max().item()
    ra.grad.copy_(total); rb.grad.copy_(total); la, lb = ra.apply(), rb.apply()
    print(it, 'grad diff', gd, 'max', gm, 'sse', total[-2].item(), one.grad[-2].item(), 'loss', float(la), float(l1), 'param diff', (ra.flat-one.flat).abs().max().item())
PY
