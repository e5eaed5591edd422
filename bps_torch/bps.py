from dex_gan_grasp_b200.bps import bps_torch  # noqa: F401
