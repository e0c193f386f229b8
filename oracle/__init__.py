"""CPU oracle for the latent-diffusion hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and there only as the checker or the
timed CPU baseline.  The product package (``stable_diffusion_pytorch_b200``)
never imports this package and has no CPU fallback.

Parity status: PINNED for the UNet forward, DDPM scheduler tables/step, CFG
combine, sampling loop and VAE decode -- ``tests/golden/make_golden.py`` ran the
unmodified reference (imported from /root/reference in the build container) on
seeded weights/inputs and committed its outputs under ``tests/golden/``;
``tests/test_oracle_golden.py`` checks this restatement against them.
DDIM is absent from the reference (no ``ddim`` anywhere in its tree), so the
DDIM step here is "parity unpinned": it restates the published eta=0 update on
the reference's own tables and x0 formula.
"""
