"""Experiment 8d: the training driver, laid out like 8d/main.py -- module-level input handles (:25-26), TrainingException
(:29-30), train(sess, batch_size, training_epochs, display_step) (:33-94), then the session block that runs it (:98-107).
The loop body is experiments/driver.py: run_8c_style."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

try:
    from . import overlap_input
    from .constants import FLAGS
    from .vae import TangoEncoder
except ImportError:                      # run as a script from inside the directory, like the reference
    import overlap_input
    from constants import FLAGS
    from vae import TangoEncoder
from experiments.driver import TrainingException, run_8c_style  # noqa: E402,F401
from magic_b200.tango import Session  # noqa: E402

# Get input data
images_batch, labels_batch = overlap_input.inputs(normalize=True, reshape=True, rotation=FLAGS.ROTATE)
n_samples = FLAGS.NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN


def train(sess, batch_size=FLAGS.BATCH_SIZE, training_epochs=60, display_step=5, **model_kw):
    vae = TangoEncoder(sess=sess, batch_size=batch_size, **model_kw)
    return run_8c_style(vae, images_batch, labels_batch, n_samples, batch_size, training_epochs, display_step)


if __name__ == "__main__":
    with Session() as sess:
        vae = train(sess=sess, training_epochs=50)
