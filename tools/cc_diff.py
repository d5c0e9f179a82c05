import numpy as np, sys
a = np.load("/tmp/cc/cc_dense.npz"); b = np.load("/tmp/cc/cc_sparse.npz")
print("forward records dense == sparse:", np.array_equal(a["best20"], b["best20"]))
n_desc = 50000
for rep in range(3):
    oa, ob = a["offsets0"], b[f"offsets{rep}"]
    bad = np.nonzero(np.diff(oa) != np.diff(ob))[0]
    print("sparse call", rep, "total", ob[-1], "dense", oa[-1], "pairs differing", bad)
    for p in bad[:3]:
        qa, ta = a["q0"][oa[p]:oa[p + 1]], a["t0"][oa[p]:oa[p + 1]]
        qb, tb = b[f"q{rep}"][ob[p]:ob[p + 1]], b[f"t{rep}"][ob[p]:ob[p + 1]]
        miss = ~np.isin(qa, qb); extra = ~np.isin(qb, qa)
        # slots of the reverse item from the forward records of this pair
        fwd = a["best20"][p * n_desc:(p + 1) * n_desc, 0]
        mark = np.zeros(n_desc, bool); mark[fwd[fwd >= 0]] = True
        slot = np.full(n_desc, -1)
        for c in range(8):
            idx = np.nonzero(mark[c::8])[0] * 8 + c
            slot[idx] = np.arange(len(idx)) * 8 + c
        tm = ta[miss]; te = tb[extra]
        print(" pair", p, "missing", miss.sum(), "extra", extra.sum())
        print("  missing t", tm[:16], "slot", slot[tm][:16], "tile", slot[tm][:16] // 128, "lane", slot[tm][:16] % 128)
        print("  extra   t", te[:16], "slot", slot[te][:16], "tile", slot[te][:16] // 128, "lane", slot[te][:16] % 128)
        print("  tiles of the item", (slot.max() // 128) + 1, "class counts", [int(mark[c::8].sum()) for c in range(8)])
