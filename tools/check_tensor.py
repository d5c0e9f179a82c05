"""GPU debug helper: tensor (tcgen05) variant of kernel 1 against the POPC variant and the oracle."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ssp_b200
from ssp_b200 import synthetic as synth
from oracle import hamming_oracle as ho

VARIANT = sys.argv[1] if len(sys.argv) > 1 else "tensor"
eng = ssp_b200.Engine(0)
ok = True
for n_img, n_desc, world in ((2, 300, 400), (3, 1500, 2000), (3, 5000, 8000)):
    sets = synth.descriptor_bank(n_img, n_desc, seed=3, world=world)
    sets[-1] = sets[-1][: n_desc - 37]          # ragged
    pairs = synth.all_pairs(n_img)
    eng.upload_descriptor_bank(sets)
    for mode in ("knn2_ratio", "crosscheck"):
        ref = eng.hamming_match_pairs(pairs, mode=mode, variant="popc", return_best2=True)
        t0 = time.time()
        got = eng.hamming_match_pairs(pairs, mode=mode, variant=VARIANT, return_best2=True)
        same = np.array_equal(ref.best2, got.best2) and np.array_equal(ref.offsets, got.offsets) and \
            np.array_equal(ref.query_idx, got.query_idx) and np.array_equal(ref.train_idx, got.train_idx) and \
            np.array_equal(ref.distance, got.distance)
        print(n_img, n_desc, mode, "same" if same else "DIFFERENT", eng.hamming_last_stats(), flush=True)
        if not same:
            ok = False
            bad = np.nonzero((ref.best2 != got.best2).any(axis=1))[0]
            print("  first differing rows:", bad[:8], "of", len(bad))
            for r in bad[:4]:
                print("   row", r, "popc", ref.best2[r], "tensor", got.best2[r])
tie = synth.tie_stress_bank(3, 700, seed=5)
eng.upload_descriptor_bank(tie)
for mode in ("knn2_ratio", "crosscheck"):
    ref = eng.hamming_match_pairs(synth.all_pairs(3), mode=mode, variant="popc", return_best2=True)
    got = eng.hamming_match_pairs(synth.all_pairs(3), mode=mode, variant=VARIANT, return_best2=True)
    same = np.array_equal(ref.best2, got.best2) and np.array_equal(ref.query_idx, got.query_idx)
    print("tie-stress", mode, "same" if same else "DIFFERENT", flush=True)
    ok &= same
print("ALL OK" if ok else "FAILED")
