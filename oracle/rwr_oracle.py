"""ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's diffusion-profile path
(multiscale/diff_prof/diffusion_profiles.py) on index arrays instead of a networkx graph.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this.  Parity status: PINNED AGAINST THE REFERENCE ITSELF — oracle/gen_golden.py runs the
reference's DiffusionProfiles (with the five networkx-3 / scipy>=1.12 shims of SURVEY §0.5) on a
synthetic MSI object and asserts bit-equal profiles and equal iteration counts; the reference's
own golden set (data/top_msi, multiscale/tests/diff_prof.py:7-45) is not in the repository.
"""
import numpy as np
import scipy.sparse as sp


def initial_matrix(n, src, dst, weight):
    """get_initial_M (:21-27): M[u,v] = w(u->v), CSR fp64, nodelist order = index order."""
    m = sp.csr_matrix((np.asarray(weight, dtype=np.float64), (src, dst)), shape=(n, n))
    m.sort_indices()
    return m


def seed_matrix(m0, seed, seed_proteins, sinks):
    """convert_M_to_make_all_drugs_indications_sinks_except_selected (:29-47): a copy of M with
    the entries protein->seed zeroed and every out-edge of every other drug/indication zeroed
    (explicit zeros stay stored)."""
    m = m0.copy()
    prot = np.asarray(seed_proteins.get(int(seed), []), dtype=np.int64)
    indptr, indices, data = m.indptr, m.indices, m.data
    for q in prot:                                             # :33-36
        lo, hi = indptr[q], indptr[q + 1]
        hit = lo + np.searchsorted(indices[lo:hi], seed)
        if hit < hi and indices[hit] == seed:
            data[hit] = 0.0
    for s in sinks:                                            # :39-46
        if s != seed:
            # the reference zeroes the entries (s, p) for p in drug_or_indication2proteins[s];
            # those are all of row s in the MSI (drugs/indications only touch their proteins)
            for p in seed_proteins.get(int(s), []):
                lo, hi = indptr[s], indptr[s + 1]
                hit = lo + np.searchsorted(indices[lo:hi], p)
                if hit < hi and indices[hit] == p:
                    data[hit] = 0.0
    return m


def refine(m):
    """refine_M_S (:49-56)."""
    s = np.array(m.sum(axis=1)).flatten()
    s[s != 0] = 1.0 / s[s != 0]
    q = sp.spdiags(s.T, 0, *m.shape, format='csr')
    return sp.csr_matrix(q * m), s


def power_iteration(m, s, seed, alpha, tol, max_iter):
    """power_iteration (:65-90) with a one-hot personalisation (:58-63,:109).  Returns (x, iterations)."""
    n = m.shape[0]
    p = np.zeros(n)
    p[seed] = 1.0
    p = p / p.sum()
    is_dangling = np.where(s == 0)[0]
    x = np.repeat(1.0 / n, n)
    for it in range(1, max_iter + 1):
        xlast = x
        x = alpha * (x * m + sum(x[is_dangling]) * p) + (1 - alpha) * p
        err = np.absolute(x - xlast).sum()
        if err < n * tol:
            return x, it
    raise RuntimeError('power iteration failed to converge in %d iterations.' % max_iter)


def diffusion_profile(m0, seed, seed_proteins, sinks, alpha, tol, max_iter):
    """calculate_diffusion_profile (:103-110) for one seed."""
    m = seed_matrix(m0, seed, seed_proteins, sinks)
    m, s = refine(m)
    return power_iteration(m, s, seed, alpha, tol, max_iter)


def clean_file_name(name):
    """:92-93."""
    return "".join([c for c in name if c.isalpha() or c.isdigit() or c == ' ' or c == "_"]).rstrip()
