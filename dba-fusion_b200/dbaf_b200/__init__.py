"""Host-side helpers around `droid_backends` for B200: the CorrBlock / AltCorrBlock mirror
(corr.py), edge-sharded global bundle adjustment (global_ba.py) and synthetic windows (synth.py)."""
