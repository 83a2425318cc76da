"""Experiment 11a: the training driver, laid out like 11a/main.py -- module-level input handles (:37-38),
train(sess, batch_size, training_epochs) (:42-133) with the joint fetch and the 6-tuple partial_fit, then the session
block that runs it.  The loop body is experiments/driver.py: run_11a_style."""
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
from experiments.driver import TrainingException, run_11a_style  # noqa: E402,F401
from magic_b200.tango import Session  # noqa: E402

# Get input data
images_batch, labels_batch = overlap_input.inputs(normalize=True, reshape=True, rotation=FLAGS.ROTATE)
n_samples = FLAGS.NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN


def train(sess, batch_size=FLAGS.BATCH_SIZE, training_epochs=60, **model_kw):
    vae = TangoEncoder(sess=sess, batch_size=batch_size, **model_kw)
    return run_11a_style(vae, sess, images_batch, labels_batch, n_samples, batch_size, training_epochs)


if __name__ == "__main__":
    with Session() as sess:
        vae = train(sess=sess, training_epochs=FLAGS.MAX_EPOCHS * 10)
