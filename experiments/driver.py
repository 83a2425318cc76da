"""The training loop every experiment directory shares (the caller of the hot path).

Mirrors `train()` of the reference drivers (8c/main.py:33-94; 11a/main.py:42-133): epochs x (n_samples / batch)
steps of `vae.partial_fit`, the NaN guard that raises TrainingException (8c/main.py:29-30,51-52), the per-step
loss line in the reference's format (8c/main.py:58-63, visible at 8c/nohup.out:27), a prediction MSE every 24
steps (8c/main.py:66-80; 11a inverts its reciprocal distances first, 11a/main.py:100), KeyboardInterrupt returning
the model (8c/main.py:91-92).  Left out on purpose: the HTTP "poor man's tensorboard" GET (8c/utils.py:250-260;
no network), the matplotlib plots (cold, not installed) and 9a's sleep(10) + PNG dumps per step (9a/main.py:50-64).

Batches come from the on-device generator (`vae.next_device_batch()`), images and labels jointly -- the fix the
reference made in 10a/main.py:64 for the mismatched double dequeue of 8c-9b."""
from __future__ import annotations

import sys
import time

import numpy as np


class TrainingException(Exception):                     # 8c/main.py:29-30, 11a/excps.py:3-4
    pass


def _line(epoch, epochs, i, total, cost, training_loss, rec_loss, lat_loss, def_loss) -> str:
    fill = " "
    return (("Epoch: (" + str(epoch) + "/" + str(epochs) + ");").ljust(16, fill)
            + (" i: (" + str(i) + "/" + str(total) + ").").ljust(16, fill) + " "
            + ("TRAINING LOSS: " + str(float(training_loss))).ljust(35, fill) + " "
            + ("Current cost: " + str(cost)).ljust(30, fill) + " "
            + ("[Reconstruction loss: " + str(rec_loss) + ";").ljust(32, fill)
            + (" latent loss: " + str(lat_loss) + ";").ljust(25, fill)
            + ("deformation loss: " + str(def_loss) + ".").ljust(27, fill) + "]")


def train(model_cls, FLAGS, sess=None, batch_size=None, training_epochs=50, display_step=5, steps_per_epoch=None,
          log=sys.stdout, seed=0, dataset=None, **model_kw):
    """`dataset`: a magic_b200.overlap_input.OverlapDataset (the PNG pairs + OVERLAP_AREAS of FLAGS.DATA_DIR, row N2);
    None trains on pairs produced by the on-device generator (row N1)."""
    batch_size = int(batch_size or FLAGS.BATCH_SIZE)
    vae = model_cls(sess=sess, batch_size=batch_size, **model_kw)
    n_samples = len(dataset) if dataset is not None else FLAGS.NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN
    total_batch = int(steps_per_epoch or max(1, n_samples // batch_size))
    data_iter = None
    if dataset is not None:
        from magic_b200.overlap_input import device_batches
        data_iter = device_batches(vae, dataset, epochs=training_epochs * 2 + 2, shuffle=True, seed=seed)
    t0, pairs = time.time(), 0
    try:
        for epoch in range(training_epochs):
            for i in range(total_batch):
                batch = next(data_iter) if data_iter is not None else vae.next_device_batch(seed)
                out = vae.partial_fit(batch, overlap_areas=batch.labels)
                cost, training_loss, rec_loss, lat_loss, def_loss = out[:5]
                pairs += batch_size
                if np.isnan(cost):
                    raise TrainingException("Got cost=nan")
                if log is not None and (i % display_step == 0 or i == total_batch - 1):
                    print(_line(epoch, training_epochs, i, total_batch, cost, training_loss, rec_loss, lat_loss, def_loss),
                          file=log)
                if i % 24 == 0:
                    test = next(data_iter) if data_iter is not None else vae.next_device_batch(seed + 1)
                    areas = test.labels if data_iter is not None else test.labels.eval()
                    predictions = vae.get_predictions(test, overlap_areas=test.labels)
                    if vae.reciprocal_distance:
                        predictions = 1.0 / predictions             # 11a/main.py:100
                    mse = float(np.mean((predictions - areas) ** 2))
                    if log is not None:
                        print("Mean squared error on a fresh batch: " + str(mse), file=log)
    except KeyboardInterrupt:
        pass
    if log is not None:
        dt = time.time() - t0
        print("trained %d pairs in %.2f s (%.0f pairs/s including logging and evaluation)" % (pairs, dt, pairs / max(dt, 1e-9)),
              file=log)
    return vae
