"""
oracle/ -- CPU restatement of the NicheCompass VGPGAE hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``nichecompass_b200/`` may import this package.  The only legal users are
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` (there only as the checker / the timed CPU baseline, never as the product path).

What it restates (all citations relative to /root/reference/src/nichecompass):
  * pyg_semantics.py -- the third-party arithmetic the reference calls but does not vendor:
    torch_geometric GATv2Conv / GCNConv / gcn_norm / softmax / self-loop utils and
    torch_sparse SparseTensor matmul (call sites nn/encoders.py:107-140,
    nn/aggregators.py:113-117,148-156,205-211,260-263).  PyG is *unpinned* in the reference
    (pyproject.toml:31) and not installable here, so these follow PyG's documented behaviour
    (SURVEY.md App. A).
  * vgpgae_ref.py    -- modules/vgpgae.py:541-1143,1146-1334,1368-1437, modules/losses.py,
    nn/encoders.py:144-208, nn/aggregators.py, nn/decoders.py, nn/layers.py,
    nn/layercomponents.py, nn/utils.py as one functional forward/loss over a state_dict.
  * sampler_ref.py   -- data/datasets.py:54-113, data/dataprocessors.py:63-95,125-129,
    data/dataloaders.py:61-102 plus PyG's loader / transform semantics.
  * train_ref.py     -- train/trainer.py:339-454 (Adam step, epoch loop) on CPU.

Pinning status: the reference ships no golden vectors / KATs for this path
(tests/test_basic.py:6-12 only checks the version string).  The oracle is therefore pinned
against *the reference's own unmodified source files executed in the builder container*
(``oracle/pyg_shim`` supplies stand-ins for the missing torch_geometric / torch_sparse /
mlflow imports; ``tests/golden/make_golden.py`` is the generating script and the vectors it
wrote are committed under ``tests/golden/``).  The PyG arithmetic itself remains
"parity unpinned": no real PyG build exists in this image to run against.
"""
