"""The stated tolerances of the path, in ONE place (DESIGN.md §2 quotes this file; tests/, __graft_entry__.smoke() and
bench.py's `parity_check` import it).  BASELINE.json north_star: BPS indices bit-exact, distances 1e-6 relative (bit-exact
is asserted), generator / evaluator outputs within 1e-4 absolute in fp32, "a bf16 tensor-core mode must reproduce the same
filtered grasp set within a stated score tolerance", MAGD within 1e-3.
"""

# fp32 modes (CUDA-core `fp32`, and the split-bf16 tensor-core `fp32_tc`): absolute, against the reference modules' outputs
FP32_TOL = 1e-4

# bf16 tensor-core mode, ONE stage on identical inputs (hidden activations and weights bf16, fp32 accumulation; BPS / z / pose
# inputs carried as hi+lo pairs): evaluator / discriminator score, generator pose outputs (rot6d, transl, joints)
BF16_SCORE_TOL = 1e-2
BF16_POSE_TOL = 2e-2

# bf16 tensor-core mode, the WHOLE pipeline as benchmarked (bf16 generator -> its poses scored by the bf16 evaluator ->
# per-object top-k): |p_bf16 - p_fp32-oracle| per grasp, the sum of the generator's pose drift seen through the evaluator and
# the evaluator's own drift.  This single epsilon also defines the filtered-set contract: per object, every row whose ORACLE
# score is more than EPS above the cut (max(thresh, k-th best oracle score)) must be kept, no row more than EPS below it may
# be; rows within EPS of the cut may fall either way.
BF16_PIPELINE_EPS = 1e-2

# bf16 tensor-core TRAINING step (C5) against the fp32 oracle step from identical state: losses relative, gradients in the
# relative L2 norm per tensor / per tensor of < 64 elements (head biases: sums of cancelling per-sample terms) / over all
# gradients of a network concatenated
BF16_TRAIN_LOSS_TOL = 2e-2
BF16_TRAIN_GRAD_TOL = 2e-1
BF16_TRAIN_GRAD_TOL_SMALL = 3.5e-1
BF16_TRAIN_GRAD_TOL_ALL = 5e-2

# MAGD on data/eval_batch.pth (north_star)
MAGD_TOL = 1e-3
