"""Drop-in mirror of multiscale/diff_prof/diffusion_profiles.py:12-169 (class DiffusionProfiles).

Same constructor, same public methods (the per-seed ones included: convert_M_to_make_all_drugs_indications_sinks_
except_selected, refine_M_S, power_iteration, with the reference's signatures and return types), same on-disk artefacts (`<clean(seed)>_p_visit_array.npy`
fp64[N] in msi.nodelist order, plus whatever msi.save_graph / save_node2idx write), but all seeds are
solved at once on the GPU by gdr_rwr_solve_f64 (csrc/rwr.cu): the per-seed matrix surgery
(:29-47), the row normalisation (:49-56) and the power iteration (:65-90) run as one batched,
bit-order-faithful fp64 kernel sequence with a per-seed convergence mask, so every seed stops at the
iteration the reference stops at.  `num_cores` is kept in the signature and ignored: the reference's
fork-per-batch fan-out (:136-155) cannot coexist with a CUDA context and is not needed.
"""
import math
import os

import numpy as np
import scipy.sparse as sp
import torch

from . import _ffi
from .graph import DeviceCSR

WEIGHT = 'weight'
_SEED_CHUNK = 4096   # seeds per device solve: 3 * N * chunk * 8 bytes of workspace (2.9 GB at N = 30k)


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("gcn-drug-repurposing_b200 needs a CUDA device: there is no CPU implementation")
    return torch.device("cuda", torch.cuda.current_device())


def sink_edge_flags(m, sink_pairs):
    """m: scipy CSR with sorted rows.  sink_pairs: int64 [k,2] of (t, p) with t a drug/indication and
    p one of drug_or_indication2proteins[t].  Returns uint8[nnz]: bit0 on stored entries (t,p), bit1
    on stored entries (p,t)  (diffusion_profiles.py:39-46 and :33-36)."""
    n = m.shape[0]
    flag = np.zeros(m.nnz, dtype=np.uint8)
    if len(sink_pairs) == 0 or m.nnz == 0:
        return flag
    rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(m.indptr))
    keys = rows * n + m.indices.astype(np.int64)          # ascending: rows sorted, indices sorted
    for bit, (a, b) in ((1, (0, 1)), (2, (1, 0))):
        want = sink_pairs[:, a] * n + sink_pairs[:, b]
        pos = np.searchsorted(keys, want)
        pos[pos >= len(keys)] = len(keys) - 1
        hit = keys[pos] == want
        flag[pos[hit]] |= bit
    return flag


def scaled_matrix_like_scipy(m, scaled_values):
    """The csr_matrix scipy's `Q * M` (Q diagonal) returns, given M (sorted rows) and the products S[row]*M.data:
    zero products dropped, every row's entries in reverse stored order (csr_matmat emits its per-row linked list
    back to front), indices therefore unsorted.  Structure only; the products come from the device."""
    n = m.shape[0]
    v = np.asarray(scaled_values)
    keep = v != 0.0
    rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(m.indptr))[keep]
    cols_r, vals_r, rows_r = m.indices[keep][::-1], v[keep][::-1], rows[::-1]
    order = np.argsort(rows_r, kind="stable")          # by row; inside a row the reversed order survives
    indptr = np.zeros(n + 1, dtype=m.indptr.dtype)
    np.cumsum(np.bincount(rows, minlength=n), out=indptr[1:])
    out = sp.csr_matrix((vals_r[order], cols_r[order].astype(m.indices.dtype), indptr), shape=m.shape)
    out.has_sorted_indices = False
    return out


class RwrProblem:
    """The weighted MSI adjacency on the device in the two orders the solver walks (rows for the row
    sums, transposed rows for x*M), with the per-entry surgery flags."""

    def __init__(self, m, sink_pairs, device=None):
        dev = device or _device()
        m = sp.csr_matrix(m, dtype=np.float64)
        m.sort_indices()
        self.n = m.shape[0]
        self.nnz = m.nnz
        self.csr = DeviceCSR.from_scipy(m, dev, dtype=torch.float64)
        flag = sink_edge_flags(m, np.asarray(sink_pairs, dtype=np.int64).reshape(-1, 2))
        self.flag = torch.from_numpy(flag).to(dev)
        self.csr_t, perm = self.csr.transpose_with_perm()
        self.flag_t = torch.empty_like(self.flag)
        L = _ffi.lib()
        _ffi.check(L.gdr_permute_values(_ffi.ptr(self.flag), _ffi.ptr(perm), self.nnz, 1, _ffi.ptr(self.flag_t),
                                        _ffi.stream()), "permute_values(flags)")

    def solve(self, seed_nodes, alpha, tol, max_iter):
        """seed_nodes: int array of node indices.  Returns (profiles fp64 CUDA tensor [S, n], iterations int32 [S])."""
        L = _ffi.lib()
        dev = self.csr.device
        seeds = torch.as_tensor(np.asarray(seed_nodes, dtype=np.int32), device=dev)
        s = int(seeds.numel())
        out = torch.empty(s, self.n, dtype=torch.float64, device=dev)
        iters = torch.empty(max(s, 1), dtype=torch.int32, device=dev)[:s]
        for lo in range(0, s, _SEED_CHUNK):
            hi = min(s, lo + _SEED_CHUNK)
            ws = _ffi.workspace(L.gdr_rwr_workspace_bytes(self.n, self.nnz, hi - lo), dev)
            c, t = self.csr, self.csr_t
            _ffi.check(L.gdr_rwr_solve_f64(
                _ffi.ptr(c.rowptr), _ffi.ptr(c.col), _ffi.ptr(c.val), _ffi.ptr(self.flag),
                _ffi.ptr(t.rowptr), _ffi.ptr(t.col), _ffi.ptr(t.val), _ffi.ptr(self.flag_t),
                self.n, self.nnz, _ffi.ptr(seeds[lo:hi]), hi - lo, float(alpha), float(tol), int(max_iter),
                _ffi.ptr(out[lo:hi]), _ffi.ptr(iters[lo:hi]), _ffi.ptr(ws), ws.numel(), _ffi.stream()),
                "rwr_solve_f64")
        return out, iters


class DiffusionProfiles():
    def __init__(self, alpha, max_iter, tol, weights, num_cores, save_load_file_path):
        self.alpha = alpha
        self.max_iter = max_iter
        self.tol = tol
        self.weights = weights
        self.num_cores = num_cores
        self.save_load_file_path = save_load_file_path

    # ---- matrix --------------------------------------------------------------------------------
    def get_initial_M(self, msi):
        """:21-27 — M[u,v] = w(u->v) in msi.nodelist order, scipy CSR fp64 (kept as `initial_M`)."""
        node2idx = {node: i for i, node in enumerate(msi.nodelist)}
        n = len(node2idx)
        if n == 0:
            assert False
        rows, cols, vals = [], [], []
        for u, v, w in msi.graph.edges(data=WEIGHT, default=1):
            if u in node2idx and v in node2idx:
                rows.append(node2idx[u]), cols.append(node2idx[v]), vals.append(w)
        m = sp.csr_matrix((np.asarray(vals, dtype=float), (np.asarray(rows, dtype=np.int64),
                                                            np.asarray(cols, dtype=np.int64))), shape=(n, n))
        m.sort_indices()
        self.initial_M = m
        self._problem = None

    def _sink_pairs(self, msi):
        pairs = []
        for t in msi.drugs_in_graph + msi.indications_in_graph:
            ti = msi.node2idx[t]
            for p in msi.drug_or_indication2proteins[t]:
                pairs.append((ti, msi.node2idx[p]))
        return np.asarray(pairs, dtype=np.int64).reshape(-1, 2)

    def _get_problem(self, msi):
        if getattr(self, "_problem", None) is None:
            self._problem = RwrProblem(self.initial_M, self._sink_pairs(msi))
        return self._problem

    def solve(self, msi, selected_drugs_and_indications):
        """All given seeds at once.  Returns (profiles ndarray [S, N] fp64, iterations ndarray [S])."""
        prob = self._get_problem(msi)
        seeds = [msi.node2idx[s] for s in selected_drugs_and_indications]
        x, it = prob.solve(seeds, self.alpha, self.tol, self.max_iter)
        return x.cpu().numpy(), it.cpu().numpy()

    # ---- per-seed API of the reference ------------------------------------------------------------
    def convert_M_to_make_all_drugs_indications_sinks_except_selected(self, msi, selected_drugs_and_indications):
        """:29-47 — a copy of `initial_M` (structure kept, explicit zeros stored) with the entries protein -> selected
        and every out-edge (t, p), p in drug_or_indication2proteins[t], of every non-selected drug / indication t
        zeroed.  Index surgery only (no arithmetic): positions come from the same per-entry flags the batched device
        solve uses (`sink_edge_flags`).  Entries the reference would assign that are not stored in M (scipy then
        inserts explicit zeros) are left absent: same matrix values, and never the case on an MSI graph, whose
        edges are stored in both directions (msi.py:64-66)."""
        m = self.initial_M.copy()
        if not m.has_sorted_indices:
            m.sort_indices()
        n = m.shape[0]
        flag = sink_edge_flags(m, self._sink_pairs(msi))
        rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(m.indptr))
        sel = np.zeros(n, dtype=bool)
        sel[[msi.node2idx[t] for t in selected_drugs_and_indications]] = True
        drop = ((flag & 1).astype(bool) & ~sel[rows]) | ((flag & 2).astype(bool) & sel[m.indices])
        m.data[drop] = 0.0
        return m

    def refine_M_S(self, M):
        """:49-56 — S = 1/rowsum(M) (0 stays 0), M <- diag(S)·M.  Row sums (numpy's reduceat order), reciprocals and
        the scaling run on the device (gdr_rwr_refine_f64); returns (scipy csr_matrix, S ndarray) like the reference.
        The returned structure is what scipy's `Q * M` leaves: zero products dropped, every row's entries in
        reverse stored order (csr_matmat emits its linked list back to front), indices therefore unsorted."""
        L = _ffi.lib()
        dev = _device()
        m = sp.csr_matrix(M, dtype=np.float64)
        if not m.has_sorted_indices:
            m = m.copy()
            m.sort_indices()
        n, nnz = m.shape[0], m.nnz
        csr = DeviceCSR.from_scipy(m, dev, dtype=torch.float64)
        s_dev = torch.empty(n, dtype=torch.float64, device=dev)
        v_dev = torch.empty(max(nnz, 1), dtype=torch.float64, device=dev)[:nnz]
        _ffi.check(L.gdr_rwr_refine_f64(_ffi.ptr(csr.rowptr), _ffi.ptr(csr.col), _ffi.ptr(csr.val), n, nnz,
                                        _ffi.ptr(s_dev), _ffi.ptr(v_dev), _ffi.stream()), "rwr_refine_f64")
        S = s_dev.cpu().numpy()
        v = v_dev.cpu().numpy()
        out = scaled_matrix_like_scipy(m, v)
        return out, S

    def power_iteration(self, M, S, nodelist, per_dict):
        """:65-90 — personalised PageRank power iteration for one personalisation vector on a caller-supplied,
        row-scaled M; returns the fp64 ndarray [N].  The iteration runs on the device (gdr_rwr_power_iteration_f64)
        in the reference's accumulation order; `self.last_iterations` holds the iteration count.  Non-convergence
        raises RuntimeError (the reference raises NetworkXError, an unimported name there: NameError in effect)."""
        import ctypes
        L = _ffi.lib()
        dev = _device()
        N = len(nodelist)
        missing = set(nodelist) - set(per_dict)
        if missing:
            raise RuntimeError('Personalization dictionary must have a value for every node. Missing nodes %s' % missing)
        p = np.array([per_dict[n] for n in nodelist], dtype=float)
        m = sp.csr_matrix(M, dtype=np.float64)
        if m.shape != (N, N):
            raise RuntimeError(f"power_iteration: M is {m.shape}, nodelist has {N} entries")
        csr_t = DeviceCSR.from_scipy(m, dev, dtype=torch.float64).transpose()      # x*M walks M.T, sources ascending
        s_dev = torch.from_numpy(np.ascontiguousarray(np.asarray(S, dtype=np.float64))).to(dev)
        p_dev = torch.from_numpy(p).to(dev)
        x = torch.empty(N, dtype=torch.float64, device=dev)
        ws = _ffi.workspace(L.gdr_rwr_power_iteration_workspace_bytes(N), dev)
        iters = ctypes.c_int32(0)
        _ffi.check(L.gdr_rwr_power_iteration_f64(_ffi.ptr(csr_t.rowptr), _ffi.ptr(csr_t.col), _ffi.ptr(csr_t.val), N,
                                                 _ffi.ptr(s_dev), _ffi.ptr(p_dev), float(self.alpha), float(self.tol),
                                                 int(self.max_iter), _ffi.ptr(x), ctypes.byref(iters), _ffi.ptr(ws),
                                                 ws.numel(), _ffi.stream()), "rwr_power_iteration_f64")
        self.last_iterations = int(iters.value)
        return x.cpu().numpy()

    def get_personalization_dictionary(self, nodes_to_start_from, nodelist):
        personalization_dict = dict.fromkeys(nodelist, 0)
        N = len(nodes_to_start_from)
        for node in nodes_to_start_from:
            personalization_dict[node] = 1. / N
        return personalization_dict

    def clean_file_name(self, file_name):
        return "".join([c for c in file_name if c.isalpha() or c.isdigit() or c == ' ' or c == "_"]).rstrip()

    def save_diffusion_profile(self, diffusion_profile, selected_drug_or_indication):
        f = os.path.join(self.save_load_file_path,
                         self.clean_file_name(selected_drug_or_indication) + "_p_visit_array.npy")
        np.save(f, diffusion_profile)

    def calculate_diffusion_profile(self, msi, selected_drugs_and_indications):
        assert (len(selected_drugs_and_indications) == 1)  # as upstream (:104)
        x, _ = self.solve(msi, selected_drugs_and_indications)
        self.save_diffusion_profile(x[0], selected_drugs_and_indications[0])

    def calculate_diffusion_profile_batch(self, msi, selected_drugs_and_indications_batch):
        seeds = [b[0] for b in selected_drugs_and_indications_batch]
        x, _ = self.solve(msi, seeds)
        for row, seed in zip(x, seeds):
            self.save_diffusion_profile(row, seed)

    def batch_list(self, list_, batch_size=None, num_cores=None):
        if batch_size == float('inf'):
            batched_list = [list_]
        else:
            if batch_size is None:
                batch_size = math.ceil(len(list_) / (float(num_cores)))
            batched_list = []
            for i in range(0, len(list_), batch_size):
                batched_list.append(list_[i:i + batch_size])
        return batched_list

    def calculate_diffusion_profiles(self, msi):
        """:123-155 — same side effects (graph + node2idx saved, graph weighted, one .npy per seed);
        the seeds are solved in one batched device call instead of forked processes."""
        msi.save_graph(self.save_load_file_path)
        msi.save_node2idx(self.save_load_file_path)
        msi.weight_graph(self.weights)
        self.get_initial_M(msi)
        computation_list = [[i] for i in msi.drugs_in_graph + msi.indications_in_graph]
        self.calculate_diffusion_profile_batch(msi, computation_list)

    def load_diffusion_profiles(self, drugs_and_indications):
        assert (not (self.save_load_file_path is None))
        drug_or_indication2diffusion_profile = dict()
        for drug_or_indication in drugs_and_indications:
            file_path = os.path.join(self.save_load_file_path,
                                     self.clean_file_name(drug_or_indication) + "_p_visit_array.npy")
            if (os.path.exists(file_path)):
                drug_or_indication2diffusion_profile[drug_or_indication] = np.load(file_path)
            else:
                print("Loading failed at " + str(drug_or_indication) + " | " + str(file_path))
        self.drug_or_indication2diffusion_profile = drug_or_indication2diffusion_profile
