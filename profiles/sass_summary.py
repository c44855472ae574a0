"""cuobjdump -sass instruction histograms of the hot kernels (run from the repo root after a build):
python profiles/sass_summary.py > profiles/sass_summary_r2.txt"""
import collections
import os
import re
import subprocess

OBJ = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "graphistician_b200", "csrc", "_obj")
KERN = {"gk_sbm.o": ["k_strip_stats<float, 2>", "k_sbm_seq2<float, 2>", "k_lp_from_stats<2>", "k_stats_fix<float, 2>"],
        "gk_eig.o": ["k_eig_seq5<8, 0>", "k_eig_pg_tile<float>", "k_eig_push<float, 8>"],
        "gk_eig_tc.o": ["k_eig_gfull_tc"],
        "gk_latent.o": ["k_latent_pair<float, 2, 0, 1>", "k_latent_pair<float, 2, 1, 1>"],
        "gk_ais.o": ["k_ais_rowcnt<4>", "k_ais_seq<3>", "k_ais_apply<3>"]}
KEEP = ("UTC", "UTMA", "LDTM", "STTM", "SYNCS", "REDUX", "LDGSTS", "MUFU", "POPC", "DFMA", "HMMA", "IMMA", "BAR", "SHFL")

print("# cuobjdump -sass instruction histograms of the hot kernels at HEAD (sm_100a); profiles/sass_summary.py\n")
for obj, names in KERN.items():
    sass = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, obj)], capture_output=True, text=True).stdout
    dem = subprocess.run(["cu++filt"], input=sass, capture_output=True, text=True).stdout
    blocks = re.split(r"\n\s*Function : ", dem)
    for nm in names:
        for b in blocks[1:]:
            head = re.sub(r"\((int|bool)\)", "", b.split("\n", 1)[0])
            if nm + "(" in head.replace("void ", ""):
                ops = collections.Counter()
                for line in b.split("\n"):
                    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
                    if m:
                        op = m.group(1)
                        ops[op if op.startswith(KEEP) else op.split(".")[0]] += 1
                print("%s\n  %d instructions: %s\n" % (nm, sum(ops.values()), ", ".join("%s %d" % kv for kv in ops.most_common(18))))
                break
        else:
            print("%s: not found in %s\n" % (nm, obj))
