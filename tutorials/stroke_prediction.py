"""Tutorial: tabular classification end to end -- CSV -> DataFrame clean-up -> Dataset -> TorchEstimator.

Covers the same ground as the reference's `tutorials/pytorch_example.ipynb` / `raytrain_example.ipynb` (a health-survey
table with string categories, missing values written as "N/A", class imbalance; a small network with one embedding per
categorical column fed by ``model(*columns)``), as a plain script:

    python tutorials/stroke_prediction.py [--rows 5000] [--epochs 5] [--trainer]

The table is generated on the fly (same columns as the public stroke-prediction survey the reference downloads:
id, gender, age, hypertension, heart_disease, ever_married, work_type, Residence_type, avg_glucose_level, bmi,
smoking_status, stroke); nothing is fetched from the network.  ``--trainer`` runs the second half with a hand-written
``TorchTrainer`` loop instead of the estimator.
"""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402

import raydp_b200 as raydp  # noqa: E402  (`import raydp` is the same package)
from raydp_b200.sql import functions as F  # noqa: E402
from raydp_b200.utils import random_split  # noqa: E402

WORK = ["children", "Govt_job", "Never_worked", "Private", "Self-employed"]
SMOKE = ["formerly smoked", "never smoked", "smokes", "Unknown"]


def write_survey_csv(path: str, rows: int, seed: int = 0) -> None:
    rng = np.random.default_rng(seed)
    age = rng.integers(1, 83, rows)
    gender = rng.choice(["Female", "Male"], rows, p=[0.59, 0.41]).astype(object)
    gender[rng.integers(0, rows)] = "Other"                       # the one outlier the clean-up removes
    hyper = (rng.random(rows) < 0.05 + age / 400).astype(int)
    heart = (rng.random(rows) < 0.02 + age / 700).astype(int)
    married = np.where((age > 22) & (rng.random(rows) < 0.8), "Yes", "No")
    work = np.where(age < 16, "children", rng.choice(WORK[1:], rows, p=[0.15, 0.01, 0.66, 0.18]))
    residence = rng.choice(["Urban", "Rural"], rows)
    glucose = np.round(rng.gamma(9.0, 12.0, rows) + 40, 2)
    bmi = np.round(rng.normal(28.5, 7.0, rows).clip(11, 70), 1).astype(object)
    bmi[rng.random(rows) < 0.04] = "N/A"                          # missing values arrive as a STRING
    smoke = rng.choice(SMOKE, rows, p=[0.17, 0.37, 0.16, 0.30])
    risk = -7.0 + age * 0.07 + hyper * 0.6 + heart * 0.7 + (glucose > 160) * 0.5
    stroke = (rng.random(rows) < 1 / (1 + np.exp(-risk))).astype(int)
    with open(path, "w") as fh:
        fh.write("id,gender,age,hypertension,heart_disease,ever_married,work_type,Residence_type,"
                 "avg_glucose_level,bmi,smoking_status,stroke\n")
        for i in range(rows):
            fh.write(f"{10000 + i},{gender[i]},{age[i]},{hyper[i]},{heart[i]},{married[i]},{work[i]},{residence[i]},"
                     f"{glucose[i]},{bmi[i]},{smoke[i]},{stroke[i]}\n")


def preprocess(df):
    """id dropped, "N/A" body-mass values replaced by the column mean, the 'Other' outlier removed, string categories
    mapped to small integers (python UDFs, as a user would write them), age bucketed."""
    df = df.drop("id")
    bmi_mean = df.filter(df.bmi != "N/A").agg(F.avg(F.col("bmi").cast("double"))).head()[0]

    @F.udf("float")
    def bmi_value(text):
        return float(bmi_mean) if text == "N/A" else float(text)

    df = df.withColumn("bmi", bmi_value(F.col("bmi"))).filter(df.gender != "Other")

    def coder(mapping):
        @F.udf("int")
        def code(value):
            return int(mapping[value])
        return code

    df = (df.withColumn("gender", coder({"Female": 0, "Male": 1})(F.col("gender")))
            .withColumn("ever_married", coder({"No": 0, "Yes": 1})(F.col("ever_married")))
            .withColumn("work_type", coder({w: i for i, w in enumerate(WORK)})(F.col("work_type")))
            .withColumn("Residence_type", coder({"Rural": 0, "Urban": 1})(F.col("Residence_type")))
            .withColumn("smoking_status", coder({s: i for i, s in enumerate(SMOKE)})(F.col("smoking_status"))))
    age_bucket = (F.when(F.col("age") < 18, 0).when(F.col("age") < 36, 1).when(F.col("age") < 56, 2)
                   .when(F.col("age") < 70, 3).otherwise(4))
    return df.withColumn("age_dis", age_bucket)


CATEGORICAL = [("gender", 2), ("hypertension", 2), ("heart_disease", 2), ("ever_married", 2), ("work_type", 5),
               ("Residence_type", 2), ("smoking_status", 4), ("age_dis", 5)]
NUMERIC = ["age", "avg_glucose_level", "bmi"]


class SurveyNet(nn.Module):
    """One small embedding per categorical column + the numeric columns -> MLP -> logit.  Called as
    ``model(*columns)`` with one ``[B, 1]`` tensor per feature column (the estimator's per-column dtype mode)."""

    def __init__(self):
        super().__init__()
        self.tables = nn.ModuleList(nn.Embedding(n, 2) for _, n in CATEGORICAL)
        self.norm = nn.BatchNorm1d(len(NUMERIC))
        self.body = nn.Sequential(nn.Linear(2 * len(CATEGORICAL) + len(NUMERIC), 64), nn.ReLU(),
                                  nn.Linear(64, 32), nn.ReLU(), nn.Linear(32, 1))

    def forward(self, *cols):
        cats = [emb(c.long().view(-1)) for emb, c in zip(self.tables, cols[:len(CATEGORICAL)])]
        nums = self.norm(torch.cat([c.float().view(-1, 1) for c in cols[len(CATEGORICAL):]], dim=1))
        return self.body(torch.cat(cats + [nums], dim=1))


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=5000)
    ap.add_argument("--epochs", type=int, default=5)
    ap.add_argument("--workers", type=int, default=1)
    ap.add_argument("--trainer", action="store_true", help="hand-written TorchTrainer loop instead of the estimator")
    args = ap.parse_args(argv)

    csv_path = os.path.join(tempfile.mkdtemp(prefix="survey_"), "survey.csv")
    write_survey_csv(csv_path, args.rows)
    spark = raydp.init_spark("stroke prediction", num_executors=1, executor_cores=1, executor_memory="500M")
    try:
        raw = spark.read.format("csv").option("header", "true").option("inferSchema", "true").load(csv_path)
        raw.show(5)
        raw.describe("age", "avg_glucose_level").show()
        print("rows with a missing bmi:", raw.filter(raw.bmi == "N/A").count())
        raw.rollup(raw.gender).count().show()                 # class / category balance, with the grand total
        raw.rollup(raw.stroke).count().show()

        data = preprocess(raw)
        train_df, test_df = random_split(data, [0.8, 0.2], 0)
        positives = train_df.filter(train_df.stroke == 1)     # oversample the rare class 4x
        for _ in range(3):
            train_df = train_df.unionByName(positives)
        features = [name for name, _ in CATEGORICAL] + NUMERIC
        dtypes = [torch.long] * len(CATEGORICAL) + [torch.float] * len(NUMERIC)
        torch.manual_seed(0)
        model = SurveyNet()

        if not args.trainer:
            from raydp_b200.torch import TorchEstimator
            est = TorchEstimator(num_workers=args.workers, model=model,
                                 optimizer=torch.optim.Adam(model.parameters(), lr=3e-3),
                                 loss=nn.BCEWithLogitsLoss(), feature_columns=features, feature_types=dtypes,
                                 label_column="stroke", label_type=torch.float, batch_size=64,
                                 num_epochs=args.epochs, metrics_name=["Accuracy"],
                                 metrics_config={"Accuracy": {"task": "binary"}})
            est.fit_on_spark(train_df, test_df)
            history = [h for h in est.results.metrics_history if "train_loss" in h]
            for h in history:
                print({k: (round(v, 4) if isinstance(v, float) else v) for k, v in h.items()
                       if k in ("epoch", "train_loss", "engine")})
            trained = est.get_model()
            losses = [h["train_loss"] for h in history]
        else:
            from raydp_b200 import train
            from raydp_b200.spark import spark_dataframe_to_ray_dataset
            from raydp_b200.train import ScalingConfig, TorchTrainer
            ds = spark_dataframe_to_ray_dataset(train_df, parallelism=args.workers)

            def loop(config):
                net = train.torch.prepare_model(SurveyNet())
                opt = torch.optim.Adam(net.parameters(), lr=3e-3)
                crit = nn.BCEWithLogitsLoss()
                shard = train.get_dataset_shard("train")
                for epoch in range(config["epochs"]):
                    total = steps = 0
                    for batch in shard.iter_torch_batches(batch_size=64):
                        cols = [batch[c].view(-1, 1) for c in config["features"]]
                        loss = crit(net(*cols).view(-1), batch["stroke"].float())
                        opt.zero_grad()
                        loss.backward()
                        opt.step()
                        total, steps = total + float(loss), steps + 1
                    train.report({"epoch": epoch, "train_loss": total / max(1, steps)})

            result = TorchTrainer(loop, train_loop_config={"epochs": args.epochs, "features": features},
                                  scaling_config=ScalingConfig(num_workers=args.workers),
                                  datasets={"train": ds}).fit()
            losses = [m["train_loss"] for m in result.metrics_history]
            for m in result.metrics_history:
                print({k: round(v, 4) if isinstance(v, float) else v for k, v in m.items()})
            trained = None
        print("train loss:", " -> ".join(f"{v:.4f}" for v in losses))
        return {"losses": losses, "model": trained, "train_rows": train_df.count(), "test_rows": test_df.count()}
    finally:
        raydp.stop_spark()


if __name__ == "__main__":
    main()
