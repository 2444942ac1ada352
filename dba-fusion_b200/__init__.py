"""dba-fusion_b200: B200-native dense-BA / correlation hot path for DBA-Fusion.

This directory is meant to sit on PYTHONPATH: it contains `droid_backends` (the drop-in
operator module) and `dbaf_b200` (CorrBlock mirror, edge-sharded global BA, synthetic windows).
"""
