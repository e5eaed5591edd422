"""DexGanGrasp.models.losses (reference: DexGanGrasp/models/losses.py:5-16,61-76) on libdexgg."""
from dex_gan_grasp_b200.losses import accuracy_evaluator, transl_rot_6D_l2_loss  # noqa: F401
