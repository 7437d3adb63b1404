"""NYC-taxi regression with ``TFEstimator`` (Keras model under ``MultiWorkerMirroredStrategy``).

Counterpart of the reference's ``examples/tensorflow_nyctaxi.py``.  TensorFlow is optional: this
framework's GPU engine is PyTorch + CUDA, and ``TFEstimator`` keeps the reference's API for code
that already has Keras models.  Without TensorFlow the script says so and exits cleanly.

    python examples/tensorflow_nyctaxi.py
"""
import sys

import _bootstrap  # noqa: F401

import raydp
from raydp import runtime as ray
from raydp.utils import random_split

from data_process import LABEL, csv_path, feature_names, nyc_taxi_preprocess

try:
    from tensorflow import keras
except ImportError:
    print("TensorFlow is not installed in this environment; TFEstimator needs it. "
          "See pytorch_nyctaxi.py for the same workflow on the PyTorch engine.")
    sys.exit(0)

from raydp.tf import TFEstimator

if __name__ == "__main__":
    ray.init(address="local", num_cpus=6)
    spark = raydp.init_spark("NYC Taxi Fare Prediction with RayDP", num_executors=1, executor_cores=1,
                             executor_memory="500M")
    frame = spark.read.format("csv").option("header", "true").option("inferSchema", "true").load(csv_path())
    spark.conf.set("spark.sql.session.timeZone", "UTC")
    frame = nyc_taxi_preprocess(frame)
    train_df, test_df = random_split(frame, [0.9, 0.1], 0)
    features = feature_names(train_df)

    inputs = keras.layers.Input(shape=(len(features),))
    x = inputs
    for width in (256, 128, 64, 16):
        x = keras.layers.BatchNormalization()(keras.layers.Dense(width, activation="relu")(x))
    model = keras.models.Model(inputs=inputs, outputs=keras.layers.Dense(1)(x))

    estimator = TFEstimator(num_workers=2, model=model, optimizer=keras.optimizers.Adam(1e-3),
                            loss=keras.losses.MeanSquaredError(), metrics=["mae"],
                            feature_columns=features, label_columns=LABEL, batch_size=256, num_epochs=10,
                            merge_feature_columns=True)
    estimator.fit_on_spark(train_df, test_df)
    print(estimator.get_model().summary())
    raydp.stop_spark()
    ray.shutdown()
