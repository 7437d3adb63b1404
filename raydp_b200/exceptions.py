"""Framework exception type.

Parity: reference ``core/raydp-main/src/main/scala/org/apache/spark/RayDPException.scala:20-23``
(``RayDPException extends SparkException``): errors raised by the Spark-on-runtime layer --
executor registration, recache-on-miss, master start-up -- carry this type so callers can tell
them from user-code failures.
"""


class RayDPException(RuntimeError):
    def __init__(self, message: str, cause: BaseException | None = None):
        super().__init__(message)
        self.__cause__ = cause
