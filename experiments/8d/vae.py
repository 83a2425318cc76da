"""Experiment 8d: the model class.  `FLAGS.BACKEND` picks the constructor; with 'b200' the class below is the
magic_b200 drop-in carrying this experiment's hard-coded hyper-parameters (8d/vae.py:49 (the only line that differs from 8c))."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

from magic_b200.tango import TangoEncoder as _B200TangoEncoder, cosine_distance  # noqa: E402,F401

try:
    from .constants import FLAGS
except ImportError:                      # run as a script from inside the directory, like the reference
    from constants import FLAGS

if FLAGS.BACKEND != "b200":
    raise ImportError("FLAGS.BACKEND = %r: this tree ships the 'b200' backend only; the 'tf' graph is the reference's "
                      "own vae.py" % (FLAGS.BACKEND,))


class TangoEncoder(_B200TangoEncoder):
    encoder_neurons = (500, 500, 500, 500)
    decoder_neurons = (500, 500)
    latent_dimensions = 200
    transfer_fct = "tanh"
    learning_rates = (1e-4, 1e-6)
    deformation_weight = 10.0
    reciprocal_distance = False
    returns_distance = False

    def __init__(self, sess=None, **kw):
        kw.setdefault("flags", FLAGS)
        super(TangoEncoder, self).__init__(sess, **kw)
