"""Shared-memory bank-conflict model of rowstream_kernel's staging buffer (pdlp_skeletons.cuh `Skew`).

8-byte products, 16 bank pairs per half-warp.  For a work block (<= 2048 nonzeros, 256 threads) it
counts the wavefronts of (a) the three coalesced passes over the products (thread t handles
k = t + 256 i) and (b) the row reduction (T threads per row, T from the row count as in the kernel),
for a pad of c slots after every 16 products, c = 0..3, plus the rejected "1 slot every 8".

    python scripts/bank_pad_sim.py
"""
import numpy as np

RNG = np.random.default_rng(0)


def log_t_for(nrows):
    lt = 0
    while lt < 5 and (2 << lt) * nrows <= 256:
        lt += 1
    return lt


def half_warp_cost(addrs):
    banks = {}
    for a in addrs:
        banks.setdefault(a % 16, set()).add(a)
    return max(len(v) for v in banks.values()) if banks else 0


def sim_block(lens, skew):
    nz, nrows = int(sum(lens)), len(lens)
    starts = np.concatenate([[0], np.cumsum(lens)]).astype(int)
    lt = log_t_for(nrows)
    t, groups = 1 << lt, 256 >> lt
    coalesced = 0
    for i in range(8):
        for hw in range(16):
            ks = [hw * 16 + lane + i * 256 for lane in range(16)]
            ks = [k for k in ks if k < nz]
            if ks:
                coalesced += half_warp_cost([skew(k) for k in ks])
    reduce_ = 0
    for base in range(0, nrows, groups):
        longest = max(lens[base:base + groups])
        for s in range((longest + t - 1) // t):
            for hw in range(16):
                addrs = []
                for lane16 in range(16):
                    tid = hw * 16 + lane16
                    rr = base + (tid >> lt)
                    if rr >= nrows:
                        continue
                    q = starts[rr] + (tid & (t - 1)) + s * t
                    if q < starts[rr + 1]:
                        addrs.append(skew(q))
                if addrs:
                    reduce_ += half_warp_cost(addrs)
    return coalesced, reduce_, nz


def make_block(gen):
    lens, total = [], 0
    while True:
        length = gen()
        if total + length > 2048 or len(lens) >= 1024:
            return lens
        lens.append(length)
        total += length


CASES = {
    "C3 CSC, 3 per column": lambda: 3,
    "C2 CSC, 5 per column": lambda: 5,
    "C5 CSC, 2 per column": lambda: 2,
    "C3 CSR conservation rows, Poisson(16)": lambda: max(1, int(RNG.poisson(16))),
    "C3 CSR capacity rows, 64": lambda: 64,
    "C2 CSR, Poisson(10)": lambda: max(0, int(RNG.poisson(10))),
    "uniform 8": lambda: 8,
    "uniform 16": lambda: 16,
}
PADS = {"c=0": lambda k: k, "c=1": lambda k: k + (k >> 4), "c=2": lambda k: k + 2 * (k >> 4), "c=3": lambda k: k + 3 * (k >> 4),
        "1 every 8": lambda k: k + (k >> 3)}


def main():
    print("wavefronts per 16 products: 3 x coalesced pass + reduction (ideal 3 + 1 = 4.0)")
    for name, gen in CASES.items():
        blocks = [make_block(gen) for _ in range(3)]
        cells = []
        for pad, f in PADS.items():
            c = r = n = 0
            for b in blocks:
                cb, rb, nb = sim_block(b, f)
                c, r, n = c + cb, r + rb, n + nb
            cells.append(f"{pad}: {(3 * c + r) / (n / 16):5.2f}")
        print(f"{name:40s} " + "  ".join(cells))


if __name__ == "__main__":
    main()
