"""P/util/engine.py: environment preparation for the JVM/Spark classpath — nothing to prepare here."""
import os


def exist_pyspark():
    try:
        import pyspark  # noqa: F401

        return True
    except ImportError:
        return False


def check_spark_source_conflict(spark_home, pyspark_path):
    pass


def compare_version(version1, version2):
    v1, v2 = [int(x) for x in version1.split(".")], [int(x) for x in version2.split(".")]
    return (v1 > v2) - (v1 < v2)


def get_bigdl_classpath():
    return os.getenv("BIGDL_CLASSPATH", "")


def is_spark_below_2_2():
    return False


def prepare_env():
    pass
