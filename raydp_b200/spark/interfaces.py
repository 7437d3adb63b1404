"""Mixin for estimators that train straight from Spark DataFrames.

Parity: reference ``python/raydp/spark/interfaces.py:23-39``.
"""
from typing import Any, NoReturn, Optional, Union

from raydp_b200.utils import convert_to_spark

DF = Any            # raydp_b200.sql.DataFrame or pandas-on-Spark DataFrame
OPTIONAL_DF = Optional[DF]


class SparkEstimatorInterface:
    def _check_and_convert(self, df):
        train_df, _ = convert_to_spark(df)
        return train_df

    def fit_on_spark(self, train_df: DF, evaluate_df: OPTIONAL_DF = None) -> NoReturn:
        """Fit from a Spark (or pandas-on-Spark) DataFrame; subclasses override."""
