import sys, torch
sys.path.insert(0, "/root/repo")
from tests.gpu_helpers import make_paged_case, run_attention
from vllm_metal_b200.metal import get_ops
ops = get_ops()
for seqs in ([(140, 300), (20, 20)], [(512, 2560)], [(300, 300), (17, 600)]):
    c = make_paged_case(seqs, 8, 2, 128, 16, "bfloat16", seed=11, sinks=True)
    a = run_attention(ops, c, kernel=3).clone()
    b = run_attention(ops, c, kernel=3).clone()
    o16 = torch.zeros(c["q"].shape, dtype=torch.float16, device="cuda")
    f = run_attention(ops, c, kernel=3, out=o16)
    d = (f.float() - a.float()).abs()
    print(seqs, "bf16 rerun equal:", torch.equal(a, b), "| fp16-out equal:", torch.equal(f.float(), a.float()), "max diff", d.max().item(),
          "n diff", int((d > 0).sum()), "of", d.numel(), "rows with diff", torch.nonzero(d.amax(dim=(1, 2)) > 0).flatten()[:10].tolist())
