"""CPU oracle for the FNO2D / TFNO2D / FourCastNet-AFNO rollout forward pass.

THIS PACKAGE IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it.  The product path (``dlwp_benchmark_b200``) never does:
it raises if the CUDA library is missing instead of falling back to anything here.

What is in here (plain PyTorch on CPU, fp32 or fp64):

* ``neuralop_fno``  -- from-spec restatement of the un-vendored ``neuralop`` 0.3.x
  ``FNO`` / ``TFNO`` that ``models/fno/fno.py:7,38-47,136-146`` instantiates.
  **Parity unpinned**: ``neuralop`` / ``tltorch`` / ``tensorly`` are absent from
  ``/root/reference`` and from this image, and the reference holds no tests or golden
  vectors, so this file is a recalled restatement cross-checked only by structure
  (parameter count, FFT identities).
* ``afno``          -- restatement of the in-repo FourCastNet ``AFNONet``
  (``models/fourcastnet/fourcastnet.py:41-361,530-543``).  **Parity pinned** against
  the reference source executed as-is (``ref_loader``; ``tests/golden/*.npz``).
* ``spectral2d``    -- restatement of the in-repo dense two-corner ``SpectralConv2d``
  and ``FourierResidualBlock`` (``models/unet/unet.py:15-69,780-836``).  Pinned the
  same way.
* ``wrappers``      -- ``FNO2DModule`` / ``TFNO2DModule`` rollout wrappers
  (``models/fno/fno.py:12-146``) with the working upstream loop form.
* ``metrics``       -- latitude-weighted RMSE / ACC (``scripts/evaluate.py:786-827``).
* ``ref_loader``    -- imports the real reference files under import stubs; usable only
  where ``/root/reference`` exists (this container, never the GPU box).
"""

from . import neuralop_fno, afno, spectral2d, wrappers, metrics  # noqa: F401
