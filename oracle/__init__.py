"""CPU oracle for the denoising hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` may be imported by the product package
(``music_style_transfer_ldm_b200``).  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs use it, and only
as the checker / the timed CPU baseline.

Parity status: PINNED for the model / sampler / loss path (``ldm_oracle``)
against outputs of the reference's own modules imported in the build
container (``tests/golden/make_golden.py`` -> ``tests/golden/*.npz``).
UNPINNED for the mel front end (``mel_oracle``): its arithmetic lives in
librosa 0.10.2.post1, which is absent and has no golden vectors in the
reference; it is cross-checked against torchaudio instead.
"""
