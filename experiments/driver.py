"""The loop bodies every experiment directory's `main.train()` shares (the caller of the hot path).

Two styles, as in the reference:
* `run_8c_style`  -- 8c/main.py:33-94 (8c, 8d, 8e, 9a): per-step loss line in the reference's format (8c/main.py:58-63,
  visible at 8c/nohup.out:27), prediction MSE every 24 steps (:66-80), per-epoch `Epoch: 0001 cost= ...` line (:86-88);
* `run_11a_style` -- 11a/main.py:42-133: joint `sess.run([images_batch, labels_batch])`, 6-tuple partial_fit, evaluation
  at `epoch % 10 == 0 and i == 3` with the reciprocal distances inverted (:94-100), per-epoch averages line (:124-126).
Both keep the NaN guard that raises TrainingException (8c/main.py:29-30,51-52) and return the model on
KeyboardInterrupt (8c/main.py:91-92).  Left out on purpose: the HTTP "poor man's tensorboard" GET (8c/utils.py:250-260;
no network), the matplotlib plots and example dumps (cold, not installed), 9a's sleep(10) per step (9a/main.py:50-64).
"""
from __future__ import annotations

import sys

import numpy as np


class TrainingException(Exception):                     # 8c/main.py:29-30, 11a/excps.py:3-4
    pass


def loss_line(epoch, epochs, i, total, cost, training_loss, rec_loss, lat_loss, def_loss) -> str:
    fill = " "
    return (("Epoch: (" + str(epoch) + "/" + str(epochs) + ");").ljust(16, fill)
            + (" i: (" + str(i) + "/" + str(total) + ").").ljust(16, fill) + " "
            + ("TRAINING LOSS: " + str(float(training_loss))).ljust(35, fill) + " "
            + ("Current cost: " + str(cost)).ljust(30, fill) + " "
            + ("[Reconstruction loss: " + str(rec_loss) + ";").ljust(32, fill)
            + (" latent loss: " + str(lat_loss) + ";").ljust(25, fill)
            + ("deformation loss: " + str(def_loss) + ".").ljust(27, fill) + "]")


def run_8c_style(vae, images_batch, labels_batch, n_samples, batch_size, training_epochs, display_step, log=None,
                 max_steps=None):
    log = log or sys.stdout
    steps = 0
    try:
        for epoch in range(training_epochs):
            avg_cost = 0.
            total_batch = int(n_samples / batch_size)
            for i in range(total_batch):
                batch_xs = images_batch.eval()
                overlap_areas = labels_batch.eval()
                cost, training_loss, rec_loss, lat_loss, def_loss = vae.partial_fit(batch_xs, overlap_areas=overlap_areas)[:5]
                if np.isnan(cost):
                    raise TrainingException("Got cost=nan")
                print(loss_line(epoch, training_epochs, i, total_batch, cost, training_loss, rec_loss, lat_loss, def_loss), file=log)
                if i % 24 == 0:
                    test_images = images_batch.eval()
                    test_areas = labels_batch.eval()
                    predictions = vae.get_predictions(test_images, overlap_areas=test_areas)
                    mse = ((predictions - np.asarray(test_areas)) ** 2).mean(axis=None)
                    print("mse: " + str(mse), file=log)
                avg_cost += cost / n_samples * batch_size
                steps += 1
                if max_steps is not None and steps >= max_steps:
                    return vae
            if epoch % display_step == 0:
                print("Epoch:", '%04d' % (epoch + 1), "cost=", "{:.9f}".format(avg_cost), file=log)
    except KeyboardInterrupt:
        pass
    return vae


def run_11a_style(vae, sess, images_batch, labels_batch, n_samples, batch_size, training_epochs, log=None, max_steps=None):
    log = log or sys.stdout
    steps = 0
    try:
        for epoch in range(training_epochs):
            avg_cost = 0.
            avg_loss = 0.
            total_batch = int(n_samples / batch_size)
            for i in range(total_batch):
                batch_xs, overlap_areas = sess.run([images_batch, labels_batch])
                cost, training_loss, rec_loss, lat_loss, def_loss, distance = vae.partial_fit(batch_xs, overlap_areas=overlap_areas)
                if np.isnan(cost):
                    raise TrainingException("Got cost=nan")
                if epoch % 10 == 0 and i == 3:
                    test_images, test_areas = sess.run([images_batch, labels_batch])
                    predictions = vae.get_predictions(test_images, overlap_areas=test_areas)
                    predictions = 1. / predictions                  # the distance is the inverse (11a/main.py:100)
                    print("Predictions: " + str(predictions), file=log)
                    print("Actual areas: " + str(np.asarray(test_areas)), file=log)
                    mse = (np.array(predictions - np.asarray(test_areas)) ** 2).mean(axis=None)
                    print("mse: " + str(mse) + "     ", file=log)
                avg_cost += cost / n_samples * batch_size
                avg_loss += training_loss / n_samples * batch_size
                steps += 1
                if max_steps is not None and steps >= max_steps:
                    return vae
            print(("Epoch: " + str(epoch)).ljust(20) + " " + ("Average cost: " + str(int(avg_cost)) + "").ljust(35) + " "
                  + ("Average loss: " + str(int(avg_loss)) + "").ljust(35), file=log)
    except KeyboardInterrupt:
        pass
    return vae
