#!/usr/bin/env python
"""Write SASS evidence for every kernel family into profiles/sass/: a mnemonic histogram per kernel (all template
instances) and the full listing of one representative instance per family."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "sass")
os.makedirs(OUT, exist_ok=True)
REPRESENTATIVE = {"conv_tc": ["conv_tc_kernelIfLi0ELi128ELi4", "conv_tc_kernelIfLi2ELi128ELi4", "conv_tc_kernelI13__nv_bfloat16Li0ELi128ELi4", "conv_tc_kernelI13__nv_bfloat16Li2ELi128ELi4"],
                  "comm": ["allreduce_sgd_shard_kernelILi8ELi4ELb1", "allreduce_shard_kernelILi0ELi8ELi8ELb1", "allreduce_oneshot_kernelILi0ELi8ELi2"],
                  "bn": ["bn_stats_kernelI13__nv", "bn_bwd_reduce_kernelIfLb1", "bn_apply_sums_kernelIfLb1ELb0"], "net_mega": ["net_fwd_bwd_kernel"], "seq": ["lstm_rec_fwd_kernel", "mel_frontend_kernel"], "loss": ["softmax_xent_kernelIf"], "optim": ["sgd_kernelILb1ELb1ELb0"], "pool": ["augment_kernel"]}
KEY = re.compile(r"FFMA|ACQBULK|PREEXIT|UTC\w*MMA\S*|LDTM\S*|STTM\S*|UTMALDG\S*|UTMASTG\S*|UBLKCP\S*|LDGSTS\S*|LDGMC\S*|SYNCS\S*|UTCBAR\S*|UTCATOM\S*|RED\.\S*|REDG\S*|ATOMG\S*|MEMBAR\S*|UCGABAR\S*|HMMA\S*|LDG\.E\.128\S*|STG\.E\.128\S*|CCTL\S*|MAPA\S*|LD\.E\S*|LDS\S*")
summary = ["# SASS evidence (cuobjdump -sass of build/*.o, sm_100a)\n",
           "Tensor-core path: `UTCHMMA` = tcgen05.mma (kind::f16 and kind::tf32 instances), `LDTM` = tcgen05.ld, `UTCBAR` = tcgen05.commit, `UTMALDG` = TMA tile load (cp.async.bulk.tensor),",
           "`LDGSTS` = cp.async, `SYNCS` = mbarrier, `UCGABAR` = cluster barrier.  NVLink path: `LDGMC.*ADD` = multimem.ld_reduce,",
           "`ST/STG.*STRONG.SYS` = peer / multicast stores, `LDG.*STRONG.SYS` = peer loads and flag polls.  No `HMMA` (legacy mma.sync) anywhere.",
           "Programmatic dependent launch: `PREEXIT` = griddepcontrol.launch_dependents, `ACQBULK` = griddepcontrol.wait.\n"]
for unit in ("conv_tc", "comm", "bn", "loss", "optim", "pool", "net_mega", "seq", "reducer", "symm"):
    obj = os.path.join(ROOT, "build", unit + ".o")
    if not os.path.exists(obj):
        continue
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    funcs = re.split(r"\n\s*Function : ", sass)[1:]
    summary.append("\n## csrc/%s.cu — %d kernels\n" % (unit, len(funcs)))
    summary.append("| kernel | SASS instructions | key mnemonics |\n|---|---:|---|")
    for f in funcs:
        name = f.split("\n", 1)[0].strip()
        dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip().replace("(anonymous namespace)::", "")
        dem = re.sub(r"\(.*", "", dem)
        ins = re.findall(r"^\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", f, flags=re.M)
        hist = collections.Counter(m for m in ins if KEY.fullmatch(m))
        keys = ", ".join("%s x%d" % (k, v) for k, v in sorted(hist.items(), key=lambda kv: -kv[1])[:9])
        summary.append("| `%s` | %d | %s |" % (dem[-90:], len(ins), keys))
        for rep in REPRESENTATIVE.get(unit, []):
            if rep in name:
                with open(os.path.join(OUT, "%s__%s.sass" % (unit, re.sub(r"[^A-Za-z0-9_]", "_", dem)[-80:])), "w") as fh:
                    fh.write("Function : " + f)
open(os.path.join(OUT, "README.md"), "w").write("\n".join(summary) + "\n")
print("\n".join(summary[:40]))
