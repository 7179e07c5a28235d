"""Developer probe: the same timing as sel_variants.py across experiment builds of the library (tools/build_variant.sh), one process
per build (CCG_LIBRARY).  usage: lib_variants.py build/variants/a.so build/variants/b.so ..."""
import os, subprocess, sys
code = r'''
import os, sys, torch
sys.path.insert(0, os.getcwd())
import cloudclean_b200 as cc
from cloudclean_b200 import synth
s = synth.config_scan(3, device="cuda"); pts = s["points"]; n = pts.shape[0]
eig = torch.empty((n, 3), device="cuda"); nrm = torch.empty((n, 4), device="cuda"); feat = torch.empty((n, 7), device="cuda")
ctx = cc.Context(0); q = []
for it in range(6):
    ctx.set_cloud(pts)
    ctx.pca_knn_features_into(20, cc.COV_NM1, 0, eig, nrm, feat); q.append(ctx.stats()["ms_query"])
print(f"{os.environ.get('CCG_LIBRARY', 'default')} minb={os.environ.get('CCG_SEL_MINB', '16')} stage={os.environ.get('CCG_SEL_STAGE', 'tma')}: query_ms min={min(q):.2f} mean={sum(q[1:])/5:.2f} checksum={int(eig.view(torch.int32).sum(dtype=torch.int64).item())}")
'''
runs = [(None, {})] + [(a, {}) for a in sys.argv[1:]] + [(None, {"CCG_SEL_MINB": "12"}), (None, {"CCG_SEL_STAGE": "ldg"}), (None, {})]
for lib, extra in runs:
    env = dict(os.environ); env.update(extra)
    if lib:
        env["CCG_LIBRARY"] = os.path.abspath(lib)
    else:
        env.pop("CCG_LIBRARY", None)
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    print(r.stdout.strip() or r.stderr[-500:], flush=True)
