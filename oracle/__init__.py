"""ORACLE -- CPU restatement of the reference's ragged message-passing path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
package; the product (gcnn_keras_b200/) never does and has no CPU fallback.

What it is: an op-for-op restatement (NumPy for integer work, torch-CPU for float work so that autograd
gives first and second derivatives) of the TensorFlow op chains in the reference, each function citing
the reference file:line it follows.  The arithmetic of the reference lives in TensorFlow (pinned
tensorflow==2.12.0 in env_linux.yml:228; >=2.9.0 in setup.py:22), which is not installable in this
environment, so the published TF semantics are restated: stable argsort, sorted segment_{sum,mean,max,min},
scatter_nd zero padding, tensor_scatter_nd_add, divide_no_nan, clip_by_value (zero gradient outside).

Parity pinning (tests/test_oracle_golden.py):
  PINNED by the reference's own golden vectors
    - GatherNodes vs NumPy indexing            test/test_gather.py:28-44
    - NodePosition -> NodeDistanceEuclidean -> BesselBasisLayer(10, 5.0)
                                                test/test_geom.py:79-128 + test/assets/bessel_basis_reference.npz
    - ACSFG2 (12 values) / ACSFG4 (48 values)  test/test_hdnnp.py:103-134
    - Gaussian expansion vs the reference's own NumPy code kgcnn/graph/adj.py:507-533
  UNPINNED (the reference has no numeric test): PoolingLocalEdges/Weighted/PoolingNodes, CosCutOff,
    EdgeDirectionNormalized, SchNetCFconv, PAiNNconv/Update, wACSF, GCN, RelationalMLP, EnergyForceModel and
    every gradient.  These are checked against fp64 autograd (gradcheck / gradgradcheck), finite-difference
    forces and invariance properties instead (tests/test_oracle_props.py).

float64 evaluation of this package is the "truth"; float32 evaluation keeps the reference's op order
(stable sort, sequential in-segment sums) and is the "TF-CPU-like" comparand.
"""
from .ragged import R  # noqa: F401
