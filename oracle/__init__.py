"""oracle/ -- TEST INFRASTRUCTURE ONLY.

CPU restatement of the STEM-GNN encoder/VQ hot path (reference:
/root/reference/STEM-GNN/model/encoder.py and model/vq.py).  It exists to *check* the
CUDA product path; it is never the thing shipped or measured.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` leg may
import it.  Nothing under `stem-gnn_b200/` imports this package.

Parity pinning status
---------------------
* VQ (`vq_ref.py`): PINNED.  The reference's own `model/vq.py` imports and runs in the
  build container (torch + einops only); `tests/golden/gen_golden.py` ran it and committed
  the input/output vectors under `tests/golden/vq_*.npz`; `tests/test_oracle_golden.py`
  checks this restatement against them.  The k-means initialisation (`kmeans_ref`) is pinned the same
  way (`tests/golden/km_*.npz`: the reference's own kmeans(), initial means fixed through `sample_fn`).
* Encoder (`propagate_ref.py`, `encoder_ref.py`): the reference `model/encoder.py` cannot be
  imported without torch_geometric 2.3.0 / torch_scatter 2.1.0 (environment.yml:292,314),
  which are absent from /root/reference and from this image.  The reference holds no tests,
  golden vectors or fixtures for this path.  The restatement follows the
  published PyG 2.3.0 / torch_scatter 2.1.0 algorithms (MessagePassing.propagate ->
  index_select -> message -> scatter_add_ / count.clamp(1); gcn_norm) and is anchored on the
  reference call sites encoder.py:82,94-97,124,193-199.  Golden vectors for it were produced
  by executing the UNMODIFIED reference encoder.py on top of `oracle/pyg_shim` (a stand-in
  for the handful of PyG symbols encoder.py:9-14 imports), so the Encoder layer loop,
  MySAGEConv.forward/message and MixtureSageLayer are the reference's own code, while the
  third-party aggregation arithmetic underneath is "parity unpinned by the reference".
* Augmentations / samplers (`augment_ref.py`, `sampler_ref.py`): torch_geometric.utils 2.3.0
  (`mask_feature`, `dropout_adj`, `negative_sampling`) and the NeighborLoader sampling algorithm
  (torch_sparse 0.6.15 `neighbor_sample`) are third-party code absent from /root/reference; restated
  from the published algorithms with the random draws passed in or replayed -- parity unpinned by the
  reference (it has no tests for its loop either).
"""
