"""Version switches.

Parity: reference ``python/raydp/versions.py:22-35`` chooses log4j1 vs log4j2 from the pyspark
version.  There is no JVM here, so the keys are kept as accepted-and-ignored configuration and
the "Spark version" is the DataFrame engine's compatibility level.
"""
SPARK_COMPAT_VERSION = "3.5.0"
RAY_COMPAT_VERSION = "2.50.0"

# names kept so user configs written for the reference parse without error
SPARK_LOG4J_VERSION = "log4j2"
SPARK_LOG4J_CONFIG_FILE_NAME_KEY = "log4j2.configurationFile"
SPARK_LOG4J_CONFIG_FILE_NAME_DEFAULT = "log4j2-default.properties"
RAY_LOG4J_VERSION = "log4j2"
RAY_LOG4J_CONFIG_FILE_NAME_KEY = "log4j2.configurationFile"
RAY_LOG4J_CONFIG_FILE_NAME_DEFAULT = "log4j2.xml"
