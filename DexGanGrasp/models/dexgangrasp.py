"""DexGanGrasp.models.dexgangrasp (reference: DexGanGrasp/models/dexgangrasp.py:28-822) on libdexgg."""
from dex_gan_grasp_b200.dexgangrasp import (DexGanGrasp, EarlyStopping, build_network, define_losses,  # noqa: F401
                                            init_net)
