"""Drop-in alias of the reference's `DexGanGrasp` package on the sm_100a kernels of libdexgg.

Put this repository's root ahead of the reference checkout on sys.path (or PYTHONPATH) and the reference's entry scripts —
dexgangrasp_offline.py, dexgangrasp_real_robot.py, eval.py, train.py — import the hot path from here unchanged:

    DexGanGrasp.models.networks      -> dex_gan_grasp_b200.networks      (ResBlock, Generator, Discriminator, DexGANGrasp, DexEvaluator)
    DexGanGrasp.models.dexgangrasp   -> dex_gan_grasp_b200.dexgangrasp   (DexGanGrasp wrapper, build_network, init_net)
    DexGanGrasp.models.losses        -> dex_gan_grasp_b200.losses
    DexGanGrasp.data.bps_encoder     -> dex_gan_grasp_b200.bps.BPSEncoder
    DexGanGrasp.utils.utils          -> the hot-path functions of dex_gan_grasp_b200.utils (over the reference's own module
                                        when it is importable, so its dataset / pose helpers stay available)

Everything this repository does NOT rebuild (config, datasets, writer, visualisation: SURVEY.md §2, out of scope) resolves to
the reference's own files: when DEXGANGRASP_REFERENCE_ROOT (default: /root/reference when it exists) names a checkout, its
DexGanGrasp/ directory is appended to this package's search path, behind the modules shipped here.
"""
import os as _os


def _reference_root():
    root = _os.environ.get("DEXGANGRASP_REFERENCE_ROOT")
    if root is None and _os.path.isdir("/root/reference/DexGanGrasp"):
        root = "/root/reference"
    return root if (root and _os.path.isdir(_os.path.join(root, "DexGanGrasp"))) else None


def _overlay(path_list, *sub):
    """Append <reference>/DexGanGrasp/<sub> to a package's __path__ (modules shipped here win, the rest falls through)."""
    root = _reference_root()
    if root:
        p = _os.path.join(root, "DexGanGrasp", *sub)
        if _os.path.isdir(p) and p not in path_list:
            path_list.append(p)


_overlay(__path__)
