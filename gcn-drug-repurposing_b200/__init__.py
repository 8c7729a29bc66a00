"""gcn-drug-repurposing_b200 — the B200-native (sm_100a) propagation path of
bowang-lab/gcn-drug-repurposing behind the reference's own Python call boundary.

Import as  ``gdr = importlib.import_module("gcn-drug-repurposing_b200")``  (the directory name
carries a hyphen) or through the repo-root alias module ``gdr_b200``.

Submodules mirror the reference files they stand in for:
    model      <- modules/model.py          (GSS_GNNLayer, ResidualGraphConvolutionalNetwork, GSS_loss)
    helper     <- helpers/helper.py         (gen_graph, combine_graph, preprocess_graph, convert_…)
    dataset    <- method/dataset.py         (DiffusionDataSet, DiffusionDataLoader)
    diffusion  <- multiscale/diff_prof/diffusion_profiles.py (DiffusionProfiles)
    optim      <- torch.optim.Adam as used at train.py:139-141 (one fused launch, same state layout)
    textio     <- np.savetxt('graph_embs.txt') / the .embs.txt reader of train.py:79-80,193 (host code, same bytes)
    trainer    <- train.py                  (main(args) restated + device-resident fast loop)
plus graph (DeviceCSR), synth (synthetic MSI / R-MAT inputs) and _ffi (ctypes binding of
csrc/libgdr.so, the C ABI in include/gdr.h).  The CUDA library is the only implementation:
nothing here falls back to CPU or to another backend.
"""
from . import _ffi  # noqa: F401
from . import synth  # noqa: F401
from . import optim  # noqa: F401
from . import textio  # noqa: F401
from . import kipf  # noqa: F401
from . import walker  # noqa: F401
from .graph import DeviceCSR, device_csr_of  # noqa: F401
from .model import GSS_GNNLayer, ResidualGraphConvolutionalNetwork, GSS_loss  # noqa: F401
from .helper import (gen_graph, combine_graph, preprocess_graph,  # noqa: F401
                     convert_sparse_matrix_to_sparse_tensor)
from .dataset import DiffusionDataSet, DiffusionDataLoader  # noqa: F401
from .diffusion import DiffusionProfiles, RwrProblem  # noqa: F401

__version__ = "0.1.0"
