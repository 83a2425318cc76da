"""Experiment 8c: hyper-parameters, attribute for attribute what 8c/constants.py holds, plus the two flags
the B200 backend adds (BACKEND, PRECISION).  Dead flags of the reference (USE_FP16, RESTORE*, LEARNING_RATE,
KEEP_PROBABILITY, TOWER_NAME) are kept so that `FLAGS.<name>` keeps resolving."""


class DISTANCE_METRICS:
    COSINE_DISTANCE = "Cosine distance"
    SQUARED_DIFFERENCE = "Squared difference"


class FLAGS:
    # ---- new: backend switch (the only edit an experiment needs)
    BACKEND = "b200"              # 'tf' = the reference's TensorFlow graph, 'b200' = magic_b200
    PRECISION = "bf16"            # 'bf16' (fp32 accumulate) or 'tf32' (bf16x3 split operands, fp32-equivalent)
    # ---- image / dataset geometry
    IMAGE_SIZE = 200
    NUM_EXAMPLES_TO_LOAD_INTO_QUEUE = 50
    NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN = 9600
    NUM_EXAMPLES_PER_EPOCH_FOR_EVAL = 4800
    ROTATE = True
    NUM_LAYERS = 3 if ROTATE else 2
    # ---- training
    BATCH_SIZE = 100
    DISTANCE_METRIC = DISTANCE_METRICS.COSINE_DISTANCE
    TRAINING_EPOCHS = 50
    # ---- dead in the reference, kept for attribute compatibility
    USE_FP16 = False
    TOWER_NAME = "tower"
    NUM_THREADS = 2
    RESTORE = False
    RESTORE_FROM = ""
    LEARNING_RATE = 1e-4
    USE_TENSORBOARD = False
    KEEP_PROBABILITY = 0.5
    RUN_INTERMEDIATE_TESTS = False
