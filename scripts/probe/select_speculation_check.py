"""dev check: the sector-parallel selection of sr_select (speculative run per sector + in-order
accept / redo) gives exactly the sequential greedy selection.  Pure Python model of both."""
import random, sys


def scan_sector(sp, ep, order, curv, seg, picked, direct, n):
    """returns (corner picks, flat picks, fwd mask, bwd mask); marks inside [sp, ep] go to `picked`"""
    fwd = bwd = 0
    cp, fp = [], []

    def mark(p):
        nonlocal fwd, bwd
        for q in range(p - 5, p + 6):
            if 0 <= q < n and seg[q] == seg[p]:
                if direct or sp <= q <= ep:
                    picked[q] = 1
                elif q > ep:
                    fwd |= 1 << (q - ep - 1)
                else:
                    bwd |= 1 << (sp - 1 - q)
    largest = 0
    for k in range(ep, sp - 1, -1):
        i = order[k]
        if not curv[i] > 0.1:
            break
        if picked[i]:
            continue
        largest += 1
        if largest <= 20:
            cp.append(i)
        else:
            break
        mark(i)
    smallest = 0
    for k in range(sp, ep + 1):
        i = order[k]
        if not curv[i] < 0.1:
            break
        if picked[i]:
            continue
        fp.append(i)
        smallest += 1
        if smallest >= 4:
            break
        mark(i)
    return cp, fp, fwd, bwd


def run(seed):
    rnd = random.Random(seed)
    n = rnd.randint(80, 400)
    S, E = 5, n - 5
    curv = [rnd.choice([0.0, 0.05, 0.2, 1.0, 3.0]) * rnd.random() for _ in range(n)]
    picked0 = [1 if rnd.random() < 0.1 else 0 for _ in range(n)]
    gap = [1 if rnd.random() < 0.05 else 0 for _ in range(n)]
    seg, c = [], 0
    for k in range(n):
        c += gap[k] if k > 0 else 0
        seg.append(c)
    bounds = [((S * (6 - j) + E * j) // 6, (S * (5 - j) + E * (j + 1)) // 6 - 1) for j in range(6)]
    order = list(range(n))
    for sp, ep in bounds:
        order[sp:ep + 1] = sorted(range(sp, ep + 1), key=lambda i: (curv[i], i))
    # sequential reference
    pa = picked0[:]
    ref = [scan_sector(sp, ep, order, curv, seg, pa, True, n)[:2] for sp, ep in bounds]
    # speculative
    pb = picked0[:]
    res = [scan_sector(sp, ep, order, curv, seg, pb, False, n) for sp, ep in bounds]
    redo = 0
    for j in range(1, 6):
        sp, ep = bounds[j]
        inc = res[j - 1][2]
        if not inc:
            continue
        fresh = [sp + b for b in range(5) if inc >> b & 1 and picked0[sp + b] == 0]
        hit = any(q in res[j][0] or q in res[j][1] for q in fresh)
        if hit:
            redo += 1
            for q in range(sp, ep + 1):
                pb[q] = picked0[q]
        for b in range(5):
            if inc >> b & 1:
                pb[sp + b] = 1
        if hit:
            res[j] = scan_sector(sp, ep, order, curv, seg, pb, False, n)
    for j in range(6):
        sp, ep = bounds[j]
        for b in range(5):
            if res[j][3] >> b & 1 and sp - 1 - b >= 0:
                pb[sp - 1 - b] = 1
            if j == 5 and res[5][2] >> b & 1 and ep + 1 + b < n:
                pb[ep + 1 + b] = 1
    assert [r[:2] for r in res] == ref, seed
    assert pa == pb, seed
    return redo


if __name__ == "__main__":
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
    redo = sum(run(s) for s in range(N))
    print(f"{N} random rings: identical picks and flags; {redo} sectors redone ({redo / (5.0 * N):.2%} of the dependent sectors)")
