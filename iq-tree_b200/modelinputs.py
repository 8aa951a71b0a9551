"""Host-side producers of the engine's INPUT CONTRACT for synthetic workloads (bench.py, large parity tests).

In the reference these inputs come from host code that stays on the host:
  * eigen-systems  ModelGTR::decomposeRateMatrix (model/modelgtr.cpp:608) -> eigensystem_sym
    (eigendecomposition.cpp:211): eval[S], evec[S][S], inv_evec[S][S] with
    evec = diag(pi^-1/2) U, inv_evec = U^T diag(pi^1/2) for the symmetrised rate matrix;
  * discrete-Gamma category rates  RateGamma::computeRates (model/rategamma.cpp:89), mean of each quantile bin;
  * the pattern matrix (Alignment, alignment.h:50) and a tree.
Nothing here runs on the timed path; numpy/scipy are fine.
"""
import random

import numpy as np


def reversible_eigensystem(rates_upper, pi):
    """rates_upper: S*(S-1)/2 exchangeabilities (row-major upper triangle), pi: S frequencies.
    Returns eval[S], evec[S*S] (row-major, evec[x][i]), inv_evec[S*S] (inv_evec[i][x]), Q normalised to 1 subst/site."""
    pi = np.asarray(pi, dtype=np.float64)
    S = pi.size
    R = np.zeros((S, S))
    R[np.triu_indices(S, 1)] = np.asarray(rates_upper, dtype=np.float64)
    R = R + R.T
    Q = R * pi[None, :]
    np.fill_diagonal(Q, 0.0)
    np.fill_diagonal(Q, -Q.sum(axis=1))
    Q /= -(pi * np.diag(Q)).sum()
    sq = np.sqrt(pi)
    A = (sq[:, None] * Q) / sq[None, :]
    A = 0.5 * (A + A.T)
    w, U = np.linalg.eigh(A)
    evec = U / sq[:, None]
    inv_evec = U.T * sq[None, :]
    return w.copy(), np.ascontiguousarray(evec).ravel(), np.ascontiguousarray(inv_evec).ravel()


def discrete_gamma(alpha, ncat):
    """Mean rate of each of ncat equiprobable Gamma(alpha, alpha) bins (Yang 1994), props = 1/ncat."""
    from scipy.special import gammainc
    from scipy.stats import gamma
    if ncat == 1:
        return np.ones(1), np.ones(1)
    cuts = gamma.ppf(np.arange(1, ncat) / ncat, alpha, scale=1.0 / alpha)
    upper = np.concatenate([gammainc(alpha + 1.0, cuts * alpha), [1.0]])
    lower = np.concatenate([[0.0], upper[:-1]])
    rates = (upper - lower) * ncat
    return rates, np.full(ncat, 1.0 / ncat)


def random_tree_newick(ntaxa, seed, names=None, mean_len=0.1, min_len=0.005):
    """Random unrooted bifurcating tree by random joining, branch lengths min_len + Exp(mean_len)."""
    rng = np.random.default_rng(seed)
    rnd = random.Random(seed)
    names = names or ["t%d" % i for i in range(ntaxa)]

    def bl():
        return float(rng.exponential(mean_len)) + min_len

    cl = list(names)
    while len(cl) > 3:
        i, j = rnd.sample(range(len(cl)), 2)
        a, b = cl[i], cl[j]
        n = "(%s:%.6f,%s:%.6f)" % (a, bl(), b, bl())
        cl = [c for k, c in enumerate(cl) if k not in (i, j)] + [n]
    return "(" + ",".join("%s:%.6f" % (c, bl()) for c in cl) + ");"


def _parse_newick(s):
    """Minimal Newick -> nested (name, length, children) for the simulator."""
    pos = [0]

    def node():
        kids = []
        if s[pos[0]] == "(":
            pos[0] += 1
            while True:
                kids.append(node())
                if s[pos[0]] == ",":
                    pos[0] += 1
                    continue
                if s[pos[0]] == ")":
                    pos[0] += 1
                    break
        j = pos[0]
        while s[j] not in ":,();":
            j += 1
        name = s[pos[0]:j]
        pos[0] = j
        ln = 0.0
        if s[pos[0]] == ":":
            j = pos[0] + 1
            while s[j] not in ",();":
                j += 1
            ln = float(s[pos[0] + 1:j])
            pos[0] = j
        return (name, ln, kids)

    return node()


def simulate_patterns(newick, names, nptn, nstates, seed, gamma_shape=0.5):
    """Evolve nptn columns down the tree with a Jukes-Cantor-type process and Gamma(gamma_shape) site rates.
    Returns uint8 states[ntaxa][nptn] in the order of `names` (each column used as one pattern)."""
    rng = np.random.default_rng(seed)
    tree = _parse_newick(newick.strip())
    rates = rng.gamma(gamma_shape, 1.0 / gamma_shape, size=nptn)
    index = {n: i for i, n in enumerate(names)}
    out = np.zeros((len(names), nptn), dtype=np.uint8)
    S = nstates
    c = S / (S - 1.0)

    def evolve(nd, state):
        name, _, kids = nd
        if not kids:
            out[index[name]] = state
            return
        for k in kids:
            p = (1.0 - np.exp(-c * k[1] * rates)) / c
            mut = rng.random(nptn) < p
            new = np.where(mut, rng.integers(0, S, size=nptn), state).astype(np.uint8)
            evolve(k, new)

    evolve(tree, rng.integers(0, S, size=nptn).astype(np.uint8))
    return out


def simulate_distinct_patterns(newick, names, nptn, nstates, seed, gamma_shape=0.5):
    """nptn DISTINCT columns in first-occurrence order: what Alignment::addPattern (alignment.cpp:674-700) turns into exactly nptn
    patterns of frequency 1, pattern p = column p.  The reference then compresses nothing away, and a GPU run that treats every column
    as one pattern evaluates the very same alignment."""
    T = len(names)
    cols = np.zeros((0, T), dtype=np.uint8)
    k = 0
    while cols.shape[0] < nptn:
        need = nptn - cols.shape[0]
        more = simulate_patterns(newick, names, int(need * 1.6) + 64, nstates, seed + 1000 * k, gamma_shape).T
        cols = np.ascontiguousarray(np.concatenate([cols, more], axis=0))
        key = cols.view(np.dtype((np.void, T))).ravel()
        _, first = np.unique(key, return_index=True)
        cols = cols[np.sort(first)]
        k += 1
    return np.ascontiguousarray(cols[:nptn].T)


def gtr_g4_inputs(seed=1, alpha=0.7, ncat=4):
    """A fixed, seeded GTR+G4 parameter set (not fitted to anything): dict ready for PhyloTree.setModel."""
    rng = np.random.default_rng(seed)
    rates = 0.5 + 2.5 * rng.random(6)
    pi = 0.15 + rng.random(4)
    pi /= pi.sum()
    ev, evec, inv = reversible_eigensystem(rates, pi)
    r, p = discrete_gamma(alpha, ncat)
    return dict(eval=ev, evec=evec, inv_evec=inv, rates=r, props=p, state_freq=pi)


def random_reversible_inputs(nstates, seed=1, alpha=0.7, ncat=4):
    """Seeded random reversible model with S states (stands in for LG / GY shapes on synthetic data)."""
    rng = np.random.default_rng(seed)
    nr = nstates * (nstates - 1) // 2
    rates = 0.05 + rng.gamma(0.8, 1.0, size=nr)
    pi = 0.2 + rng.random(nstates)
    pi /= pi.sum()
    ev, evec, inv = reversible_eigensystem(rates, pi)
    r, p = discrete_gamma(alpha, ncat)
    return dict(eval=ev, evec=evec, inv_evec=inv, rates=r, props=p, state_freq=pi)
