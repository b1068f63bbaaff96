"""Per-read EM of the reference in plain Python floats (tests only): the same operations in the same order as
src/Algorithm.cpp:69-130 (and oracle/basetype_oracle.c o_single_em / o_em), so every value -- in particular each pass's
delta = sum_i |log ml'_i - log ml_i| -- is the IEEE double the reference computes (math.log is glibc's log).  Used to put
em_epsilon EXACTLY on a delta the reference meets (tests of BVC_FLAG_EM_DELTA); small sites only."""
import math

MLN10TO10 = -0.23025850929940458


def likelihood_rows(bases, quals):
    rows = []
    for b, q in zip(bases, quals):                       # src/BaseType.cpp:9-20
        e = math.exp(MLN10TO10 * q)
        rows.append([1.0 - e if j == b else e / 3.0 for j in range(4)])
    return rows


def single_em(f, L):
    n = len(L)
    ml = [0.0] * n
    post = [[0.0] * n for _ in range(4)]
    for i, row in enumerate(L):
        lik = [0.0] * 4
        for j in range(4):
            lik[j] = f[j] * row[j]
            ml[i] += lik[j]
        for j in range(4):
            post[j][i] = lik[j] / ml[i]
    eap = [0.0] * 4
    for j in range(4):
        for i in range(n):
            eap[j] += post[j][i]
        eap[j] = eap[j] / n
    return ml, eap


def em_deltas(bases, quals, subset, iters=100, eps=0.001):
    """The deltas of the passes the reference runs for the model `subset` (list of base codes), in order."""
    L = likelihood_rows(bases, quals)
    depth = [sum(1 for b in bases if b == j) for j in range(4)]
    dsum = sum(depth[j] for j in subset)
    f = [depth[j] / dsum if j in subset else 0.0 for j in range(4)]
    ml, eap = single_em(f, L)
    deltas = []
    for _ in range(iters):
        f = eap
        ml2, eap = single_em(f, L)
        delta = 0.0
        for a, b in zip(ml2, ml):
            delta += abs(math.log(a) - math.log(b))
        ml = ml2
        deltas.append(delta)
        if delta < eps:
            break
    return deltas
