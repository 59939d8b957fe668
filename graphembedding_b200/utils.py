"""Helpers with the reference's names (src/utils.py, model/Siamese/utils_siamese.py)."""
from .data import load_data  # noqa: F401  (src/utils.py:15-35)


def get_train_str(train):
    return 'train' if train else 'test'


def check_flags():
    """model/Siamese/utils_siamese.py:13-18."""
    from .config import FLAGS
    assert 0 < FLAGS.valid_percentage < 1
    assert FLAGS.sample_num >= -1
    assert FLAGS.yeta >= 0
    assert FLAGS.num_layers >= 2
    assert FLAGS.batch_size >= 1


def print_msec(sec):
    return '{:.2f}msec'.format(sec * 1000)


def get_ts():
    """src/utils.py:142-147: timestamp used in log / result file names."""
    import datetime
    return datetime.datetime.now().strftime('%Y-%m-%dT%H:%M:%S')
