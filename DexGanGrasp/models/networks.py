"""DexGanGrasp.models.networks (reference: DexGanGrasp/models/networks.py:12-337) on libdexgg."""
from dex_gan_grasp_b200.networks import (DexEvaluator, DexGANGrasp, Discriminator, Generator, ResBlock,  # noqa: F401
                                         get_scheduler)
