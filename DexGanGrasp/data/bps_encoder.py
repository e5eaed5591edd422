"""DexGanGrasp.data.bps_encoder (reference: DexGanGrasp/data/bps_encoder.py:12-55) on the sm_100a BPS kernel."""
from dex_gan_grasp_b200.bps import BPSEncoder  # noqa: F401
