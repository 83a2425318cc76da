"""Experiment 8d: the training driver (structure of 8d/main.py: train() -> epochs x batches of partial_fit,
NaN guard, per-step loss line, periodic prediction MSE).  Inputs come from the on-device generator."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

try:
    from .constants import FLAGS
    from .vae import TangoEncoder
except ImportError:
    from constants import FLAGS
    from vae import TangoEncoder
from experiments.driver import train  # noqa: E402

if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=FLAGS.TRAINING_EPOCHS)
    ap.add_argument("--batch", type=int, default=FLAGS.BATCH_SIZE)
    ap.add_argument("--steps-per-epoch", type=int, default=0)
    a = ap.parse_args()
    train(TangoEncoder, FLAGS, sess=None, batch_size=a.batch, training_epochs=a.epochs,
          steps_per_epoch=a.steps_per_epoch or None)
