"""Bring-up probe: runs one GEMM case per subprocess (so a trap in one case cannot
poison the CUDA context of the next) and prints the error against torch fp32."""
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

CASES = {
    # name: (M, N, K, a_mn, b_mn, block_n)
    "kk_128": (128, 128, 64, False, False, 128),
    "kk_k256": (128, 128, 256, False, False, 128),
    "kk_big": (512, 384, 320, False, False, 128),
    "kk_bn64": (256, 192, 128, False, False, 64),
    "kk_bn256": (256, 512, 128, False, False, 256),
    "kmn_128": (128, 128, 64, False, True, 128),
    "kmn_big": (512, 384, 320, False, True, 128),
    "kmn_bn256": (384, 512, 192, False, True, 256),
    "mnmn_128": (128, 128, 64, True, True, 128),
    "mnmn_big": (384, 256, 1000, True, True, 128),
    "mnk_big": (384, 256, 328, True, False, 64),
    "tail": (200, 588, 72, False, True, 0),
    "persist": (4096, 2304, 768, False, True, 0),
    # CTA-pair kernel (JM_GEMM_PAIR=1 in the environment forces it)
    "pair_kk_256": (256, 256, 64, False, False, 256),
    "pair_kk_k512": (256, 256, 512, False, False, 256),
    "pair_kmn_256": (256, 256, 128, False, True, 256),
    "pair_mnmn_256": (256, 256, 128, True, True, 256),
    "pair_kk_big": (1024, 768, 576, False, False, 256),
    "pair_kmn_big": (1280, 512, 512, False, True, 256),
    "pair_mnmn_big": (768, 1024, 3000, True, True, 256),
    "pair_tail": (600, 520, 200, False, True, 256),
    "pair_persist": (26624, 2304, 768, False, True, 256),
}


def run_case(name):
    import torch
    from jumbo_mae_tpu_b200 import ops
    M, N, K, a_mn, b_mn, bn = CASES[name]
    g = torch.Generator(device="cuda").manual_seed(1)
    A = torch.randn((K, M) if a_mn else (M, K), generator=g, device="cuda").to(torch.bfloat16)
    B = torch.randn((K, N) if b_mn else (N, K), generator=g, device="cuda").to(torch.bfloat16)
    Af = (A.float().t() if a_mn else A.float())
    Bf = (B.float() if b_mn else B.float().t())
    ref = Af @ Bf
    out = torch.full((M, N), float("nan"), device="cuda")
    ops.gemm(A, B, M, N, K, a_mn=a_mn, b_mn=b_mn, out_f32=out, block_n=bn)
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item()
    rel = ((out - ref).norm() / ref.norm()).item()
    print(f"{name}: max_abs_err={err:.4g} rel_l2={rel:.3g} nan={torch.isnan(out).sum().item()}", flush=True)
    if not (rel < 1e-3):
        bad = (out - ref).abs() > 0.05 * ref.abs().max()
        rows = bad.any(1).nonzero().flatten()[:8].tolist()
        cols = bad.any(0).nonzero().flatten()[:8].tolist()
        print(f"   bad rows {rows} cols {cols} frac={bad.float().mean().item():.3f}")
        print("   out[0,:8]", out[0, :8].tolist())
        print("   ref[0,:8]", ref[0, :8].tolist())
        return 1
    return 0


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] != "all":
        sys.exit(run_case(sys.argv[1]))
    fails = 0
    only = os.environ.get("PROBE_PREFIX", "")
    for name in CASES:
        if not name.startswith(only):
            continue
        env = dict(os.environ)
        if name.startswith("pair"):
            env["JM_GEMM_PAIR"] = "1"
        try:
            r = subprocess.run([sys.executable, __file__, name], timeout=120, capture_output=True, text=True, env=env)
            sys.stdout.write(r.stdout)
            if r.returncode != 0:
                fails += 1
                sys.stdout.write(f"{name}: FAILED rc={r.returncode}\n{r.stderr[-1500:]}\n")
        except subprocess.TimeoutExpired:
            fails += 1
            print(f"{name}: TIMEOUT")
        sys.stdout.flush()
    print("failures:", fails)
