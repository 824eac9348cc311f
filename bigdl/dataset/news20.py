from bigdl_b200.dataset.news20 import *  # noqa: F401,F403

from bigdl.dataset import base as _base  # noqa: E402

NEWS20_URL = "http://qwone.com/~jason/20Newsgroups/20news-19997.tar.gz"
GLOVE_URL = "http://nlp.stanford.edu/data/glove.6B.zip"


def download_news20(dest_dir):
    import tarfile
    path = _base.maybe_download("20news-19997.tar.gz", dest_dir, NEWS20_URL)
    with tarfile.open(path, "r:gz") as tar:
        tar.extractall(dest_dir)
    return dest_dir


def download_glove_w2v(dest_dir):
    import zipfile
    path = _base.maybe_download("glove.6B.zip", dest_dir, GLOVE_URL)
    with zipfile.ZipFile(path, "r") as z:
        z.extractall(dest_dir)
    return dest_dir
