"""Static SASS mnemonic counts per kernel family of the in-tree library (profiles/sass_summary.md)."""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "jumbo_mae_tpu_b200/libjumbo_mae_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
fams = [("attn_tc_fwd", "attn_tc_fwd"), ("attn_tc_bwd", "attn_tc_bwd"), ("attn_fwd_kernel", "attn_fwd"),
        ("attn_bwd_kernel", "attn_bwd"), ("gemm2_kernel", "gemm2_kernel"), ("gemm_kernel", "gemm_kernel"),
        ("layernorm", "layernorm"), ("lamb", "lamb"), ("adamw", "adamw"), ("softmax_ce", "softmax_ce"),
        ("patch_loss", "patch_loss"), ("patchify / targets", "patchify"), ("patchify / targets", "masked_targets"),
        ("rrc (augmentation)", "rrc_"), ("mask_sort", "mask_sort")]
cols = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMALDG.MULTICAST", "UTMASTG", "UTMAREDG", "HMMA", "MUFU.EX2", "MUFU.TANH"]
cnt = collections.defaultdict(lambda: collections.Counter())
kernels = collections.Counter()
fam = None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        name = m.group(1)
        fam = next((f for f, key in fams if key in name), "other")
        kernels[fam] += 1
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and fam:
        op = m.group(1)
        for c in cols:
            if c == "UTCHMMA.2CTA":
                hit = op.startswith("UTCHMMA") and "2CTA" in op
            elif c == "UTMALDG.MULTICAST":
                hit = op.startswith("UTMALDG") and "MULTICAST" in op
            else:
                hit = op.startswith(c)
            if hit:
                cnt[fam][c] += 1
print("| family | kernels | " + " | ".join(cols) + " |")
print("|---|---|" + "---|" * len(cols))
for f in kernels:
    print(f"| {f} | {kernels[f]} | " + " | ".join(str(cnt[f][c]) for c in cols) + " |")
