"""train.py of the reference (train.py:72-193) restated on the drop-in classes.

    main(args)        the reference driver, line for line, minus what cannot run on current numpy
                      (np.float, :80) — same flags (parse_args mirrors train.py:12-59), same order of
                      RNG draws, same `graph_embs.txt` output (np.savetxt, '%.18e', headerless).
    GCNTrainer        the loop body (train.py:151-187) as an object, for benchmarks and tests: one
                      `iteration(batch)` = full-graph forward + batch GSS loss + backward + Adam.

Everything numeric goes through gcn-drug-repurposing_b200's classes, i.e. libgdr.so; there is no CPU path.
"""
import argparse
import time

import numpy as np
import torch

from .dataset import DiffusionDataLoader, DiffusionDataSet
from .optim import Adam
from .textio import read_node2vec_embs, savetxt_e18
from .helper import (combine_graph, convert_sparse_matrix_to_sparse_tensor, gen_graph, gram_percentile,
                     preprocess_graph)
from .model import GSS_loss, ResidualGraphConvolutionalNetwork


def build_parser():
    """train.py:12-58, flag for flag with the reference's defaults.  `--gpu-id None` means CPU upstream
    (train.py:68); this path has no CPU implementation, so None selects the current CUDA device.  `--dataset`,
    `--data-path` and `--report-hard` are accepted and unused (they only feed the ransac graph modes, whose
    datasets are absent upstream).  `--out-file` is an addition: train.py:193 hard-codes `graph_embs.txt`."""
    p = argparse.ArgumentParser(description='GCN similarity-loss embedding on a B200')
    p.add_argument('--kq', type=int, default=5)
    p.add_argument('--k', type=int, default=5)
    p.add_argument('--alpha', type=float, default=1)
    p.add_argument('--beta', type=float, default=None)
    p.add_argument('--beta-percentile', type=float, default=None)
    p.add_argument('--seed', type=int, default=None)
    p.add_argument('--epochs', type=int, default=200)
    p.add_argument('--batch-size', type=int, default=0)
    p.add_argument('--hidden-units', type=int, default=128)
    p.add_argument('--num-layers', type=int, default=2)
    p.add_argument('--loss', type=str, default='gss')
    p.add_argument('--lr', type=float, default=0.0001)
    p.add_argument('--init-weights', type=float, default=1e-5)
    p.add_argument('--regularizer-scale', type=float, default=1e-5)
    p.add_argument('--layer-decay', type=float, default=0.3)
    p.add_argument('--dataset', type=str, default='roxford5k', choices=['roxford5k', 'rparis6k', 'instre'])
    p.add_argument('--emb-file', type=str, default=None, help='node2vec-format .embs.txt (train.py:79)')
    p.add_argument('--data-path', type=str, default=None)
    p.add_argument('--gpu-id', type=int, default=None)
    p.add_argument('--report-hard', action='store_true')
    p.add_argument('--graph-mode', type=str, default='descriptor', choices=['descriptor', 'ransac', 'approx_ransac'])
    p.add_argument('--out-file', type=str, default='graph_embs.txt', help="train.py:193 hard-codes this name")
    return p


def load_embs(path):
    """train.py:79-82: skip the 'N dim' header, drop the node-name column, fp64.  Parsed by libgdr's host-side
    reader (csrc/textio.cu) instead of np.loadtxt(dtype=object): same values, seconds instead of minutes at MSI scale."""
    _names, vals = read_node2vec_embs(path)
    return vals


class GCNTrainer:
    """Model + optimizer + loss of train.py:111-150, features and adjacency resident on the device."""

    def __init__(self, features, adj, hidden_units, num_layers=2, layer_decay=0.3, init_weights=1e-5, alpha=1.0,
                 lr=1e-4, batch_size=2048, beta=None, beta_percentile=None, device="cuda", hoist_layer1=True):
        self.device = torch.device(device)
        self.dataset = DiffusionDataSet(features=features, adj=adj.to(self.device))
        n = len(self.dataset)
        self.batch_size = batch_size if batch_size > 0 else n
        self.model = ResidualGraphConvolutionalNetwork(train_batch_size=self.batch_size, val_batch_size=n,
                                                       num_layers=num_layers, hidden_units=hidden_units,
                                                       init_weights=init_weights, layer_decay=layer_decay)
        self.model.to(self.device)
        self.model.train()
        self.optimizer = Adam(self.model.parameters(), lr=lr, weight_decay=0)       # train.py:139-141 (lr: --lr,
        #                                                                             default 1e-4; train.sh passes 3e-4)
        self.loss_fcn = GSS_loss(alpha).gss_loss
        self.beta = beta
        self.beta_percentile = beta_percentile
        self.features_dev = self.dataset.features.to(self.device)
        self.hidden_emb = None
        # The trainer owns `features_dev` and never writes it, so it may let the model reuse the first layer's
        # aggregations across iterations (model.HOIST_LAYER1).  Only for that resident tensor: features passed to
        # iteration() explicitly are somebody else's and are always aggregated afresh.
        self.hoist_layer1 = bool(hoist_layer1)

    def invalidate(self):
        """Call after changing `features_dev` or the adjacency values in place by means torch cannot see (`.data`,
        libgdr `out=` writes): drops the cached first-layer aggregations."""
        self.model.invalidate_layer1_cache()

    def loader(self, shuffle=True):
        return DiffusionDataLoader(self.dataset, batch_size=self.batch_size, num_workers=6, shuffle=shuffle)

    def select_beta(self, hidden_emb):
        """train.py:163-172."""
        if self.beta_percentile is not None:
            return gram_percentile(hidden_emb.detach(), self.beta_percentile)
        if self.beta is not None:
            return self.beta
        raise Exception('At least one of beta and beta_percentile should be set!')

    def iteration(self, batch, features=None):
        """train.py:157-184.  batch: int64 index tensor on the device."""
        x = self.features_dev if features is None else features
        self.model.hoist_layer1 = self.hoist_layer1 and features is None
        hidden_emb = self.model(x=x, adj=self.dataset.adj)
        if self.beta is None:
            self.beta = self.select_beta(hidden_emb)
        loss = self.loss_fcn(embs=hidden_emb, beta=self.beta, index=batch.detach())
        self.optimizer.zero_grad()
        loss.backward()
        self.optimizer.step()
        self.hidden_emb = hidden_emb
        return loss

    def epoch(self, batches):
        loss = None
        for b in batches:
            loss = self.iteration(b)
        return loss

    def epoch_from_host(self, batches, features_host):
        """One epoch of train.py:155-184 with HOST inputs, moved as the reference moves them — the index batch AND the
        whole feature matrix go to the device in every iteration (train.py:157-160) and the loss is read back — but
        with the upload of iteration i+1 running on a copy stream, into the other of two device buffers, while
        iteration i computes.  batches: list of int64 CPU index tensors (pinned for asynchronous copies);
        features_host: [N, d] fp32 CPU tensor (pinned).  Returns the list of loss values."""
        dev = self.device
        cur = torch.cuda.current_stream(dev)
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
            self._feat_bufs = [torch.empty(features_host.shape, dtype=torch.float32, device=dev) for _ in range(2)]
            self._buf_free = [None, None]
        cs = self._copy_stream

        def upload(i):
            with torch.cuda.stream(cs):
                if self._buf_free[i % 2] is not None:
                    cs.wait_event(self._buf_free[i % 2])       # the iteration that read this buffer has finished
                else:
                    cs.wait_stream(cur)
                self._feat_bufs[i % 2].copy_(features_host, non_blocking=True)
                bd = batches[i].to(dev, non_blocking=True)
                return bd, cs.record_event()

        losses = []
        nxt = upload(0) if batches else None
        for i in range(len(batches)):
            bd, ev = nxt
            if i + 1 < len(batches):
                nxt = upload(i + 1)
            cur.wait_event(ev)
            bd.record_stream(cur)
            loss = self.iteration(bd, features=self._feat_bufs[i % 2])
            self._buf_free[i % 2] = cur.record_event()
            losses.append(loss.item())                         # device -> host read of the step's result
        return losses


def main(args):
    """train.py:72-193."""
    if args.seed:
        torch.manual_seed(args.seed)
        np.random.seed(args.seed)
    if args.beta is not None and args.beta_percentile is not None:
        raise Exception('beta and beta_percentile can not be used at the same time!')
    if args.graph_mode != 'descriptor':
        raise Exception('ransac graph modes need the retrieval datasets, absent from the reference repository too')
    if args.emb_file is None:
        raise Exception('--emb-file is required (train.py:79 reads it unconditionally)')
    ppi_embs = load_embs(args.emb_file)
    if args.hidden_units != ppi_embs.shape[1]:
        # upstream nn.Linear(h, h) raises a shape error on the first forward (model.py:142,165)
        raise RuntimeError(f"--hidden-units {args.hidden_units} must equal the embedding width {ppi_embs.shape[1]} "
                           "(the layer is Linear(h, h) applied to adj @ features)")
    X = ppi_embs.T                                   # (d, num_nodes)
    Q = X[:, 0:3]
    q_adj, q_features, x_adj, x_features = gen_graph(Q, X, args.kq, args.k, None, None)
    x_adj_normed = preprocess_graph(x_adj)
    x_adj_normed_sparse_tensor = convert_sparse_matrix_to_sparse_tensor(x_adj_normed)
    # train.py:96-104 also builds the (N+3)-node query graph; nothing consumes it afterwards
    _ = combine_graph(q_adj, x_adj)
    trainer = GCNTrainer(x_features, x_adj_normed_sparse_tensor, args.hidden_units, num_layers=args.num_layers,
                         layer_decay=args.layer_decay, init_weights=args.init_weights, alpha=args.alpha, lr=args.lr,
                         batch_size=args.batch_size, beta=args.beta, beta_percentile=args.beta_percentile,
                         device="cuda" if args.gpu_id is None else f"cuda:{args.gpu_id}")
    if args.loss != 'gss':
        raise Exception("only the 'gss' loss is functional upstream (tri_loss references undefined names)")
    loader = trainer.loader(shuffle=True)
    itr = 0
    while itr < args.epochs:
        start_time = time.time()
        for batch_id, batch_data in enumerate(loader):
            trainer.iteration(batch_data.to(trainer.device))
        train_time = time.time() - start_time  # noqa: F841  (computed and unused upstream as well)
        itr += 1
        print(f"iter {itr}")
    # train.py:193 np.savetxt(..., hidden_emb): same bytes, rows formatted in parallel by libgdr's host code
    savetxt_e18(args.out_file, trainer.hidden_emb.detach().cpu().numpy())
    return trainer


if __name__ == '__main__':
    main(build_parser().parse_args())
