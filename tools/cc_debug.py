"""Debug aid: cross-check of a synthetic bank, repeated calls, saved for a dense-vs-sparse diff."""
import os, sys, numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import ssp_b200
from ssp_b200 import synthetic as synth

n_img, n_desc, out = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
sets = synth.descriptor_bank(n_img, n_desc, seed=0)
pairs = synth.all_pairs(n_img)
eng = ssp_b200.Engine(0)
eng.upload_descriptor_bank(sets)
runs = []
for rep in range(3):
    res = eng.hamming_match_pairs(pairs, mode="crosscheck", return_best2=True)
    runs.append((np.asarray(res.offsets).copy(), res.query_idx.copy(), res.train_idx.copy(), res.best2.copy()))
    print("call", rep, "survivors", int(res.offsets[-1]), "best2 equal to call 0:", np.array_equal(runs[-1][3], runs[0][3]), flush=True)
np.savez(out, **{f"{k}{i}": v for i, r in enumerate(runs) for k, v in zip(("offsets", "q", "t", "best2"), r)})
