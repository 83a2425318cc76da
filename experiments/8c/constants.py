"""Experiment 8c: hyper-parameters -- every attribute of 8c/constants.py:12-89 under its own name and value, plus the
two flags the B200 backend adds (BACKEND, PRECISION).  Dead flags of the reference (USE_FP16, RESTORE*, LEARNING_RATE,
KEEP_PROBABILITY, TOWER_NAME, the e-mail / download / checkpoint settings) keep resolving."""


class DISTANCE_METRICS:
    COSINE_DISTANCE = "Cosine distance"
    SQUARED_DIFFERENCE = "Squared difference"


class FLAGS:
    def __init__(self):
        pass

    # ---- new: backend switch (the only edit an experiment needs)
    BACKEND = "b200"              # 'tf' = the reference's TensorFlow graph, 'b200' = magic_b200
    PRECISION = "bf16"            # 'bf16' (fp32 accumulate) or 'tf32' (bf16x3 split operands, fp32-equivalent)

    # ---- dataset info: PNG pairs '%d_L.png' / '%d_K.png' + OVERLAP_AREAS under DATA_DIR; when the directory does not
    # exist the B200 backend's inputs() falls back to the on-device pair generator
    DATA_DIR = '/data/affinity/2d/overlap_tiny'
    DATA_URL = 'http://nowhere'
    INTERNET = True
    ALLOW_DOWNLOADS = False
    IMAGE_SIZE = 200

    # ---- training info
    NUM_EXAMPLES_TO_LOAD_INTO_QUEUE = 50
    NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN = 9600
    NUM_EXAMPLES_PER_EPOCH_FOR_EVAL = 4800
    MAX_EPOCHS = 10
    TRAINING_STEP = 10000
    MAX_TRAIN_ITERATIONS = 10000
    ROTATE = True
    NUM_LAYERS = 3 if ROTATE else 2
    BATCH_SIZE = 100
    DISTANCE_METRIC = DISTANCE_METRICS.COSINE_DISTANCE
    USE_FP16 = False
    TOWER_NAME = 'tower'
    NUM_THREADS = 2
    NOTIFICATION_EMAIL = ''       # the reference stores its author's address here; nothing is ever sent
    EMAIL_INFO = False
    RESTORE = False
    RESTORE_FROM = '../../../test_data/summaries/netstate/saved_state-91000'
    CHECKPOINT_DIR = './summaries'
    LEARNING_RATE = 1e-4
    CHECK_DATASET = True
    TRAIN_DIR = './train'
    SUMMARIES_DIR = './summaries'
    USE_TENSORBOARD = False
    KEEP_PROBABILITY = 0.5
    RUN_NAME = 'test'
    PRINT_INFO = True
    SEND_LOSS_TO_POOR_MANS_TENSORBOARD_EVERY_N_STEPS = 50
    SAVE_NETSTATE_AND_EMAIL_STATUS_EVERY_N_STEPS = 1000
