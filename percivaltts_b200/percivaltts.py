"""
Configuration object and logging helpers used by the callers of the hot path.
Mirrors the names of reference percivaltts/percivaltts.py:51-142 (readids, configuration, print_log, print_tty,
time2str, is_int); plots and system probes of that module are out of scope.
"""
import datetime
import runpy
import sys

import numpy as np


def numpy_force_random_seed(seed=123):
    """reference percivaltts.py:30-33: one global numpy seed for repeatable shuffles / crops."""
    np.random.seed(seed)


def readids(fileids):
    """File ids, one per line, blank lines dropped (percivaltts.py:51-54)."""
    with open(fileids) as f:
        return [ln.strip() for ln in f if ln.strip()]


class configuration(object):
    """Plain attribute bag carried along the pipeline (percivaltts.py:57-104)."""

    def __eq__(self, other):
        return self.__dict__ == other.__dict__

    def __ne__(self, other):
        return not self.__eq__(other)

    def id_train_nb(self):
        return self.train_batch_size * int(np.floor(self.id_valid_start / self.train_batch_size))

    def _public_items(self):
        return sorted((k, v) for k, v in self.__dict__.items() if not k.startswith('__') and not callable(v))

    def print_content(self):
        hypers = set(h[0] for h in getattr(self, 'train_hypers', []) or [])
        for key, val in self._public_items():
            note = '    (Attention! hyper-parameter optimized during multi-training)' if key in hypers else ''
            print('    {:<30}{}{}'.format(key, val, note))
        print('')

    def mergefiles(self, filenames):
        if not isinstance(filenames, (list, tuple)):
            filenames = [filenames]
        for fname in filenames:
            for k, v in runpy.run_path(fname).items():
                setattr(self, k, v)

    def merge(self, cfgtoadd):
        for k, v in cfgtoadd.__dict__.items():
            if not k.startswith('__'):
                setattr(self, k, v)


def datetime2str(sec=-1):
    d = datetime.datetime.now() if sec < 0 else datetime.datetime.fromtimestamp(sec)
    return d.strftime('%Y-%m-%d %H:%M:%S')


def time2str(sec):
    sec = float(sec)
    days, rem = divmod(int(sec), 86400)
    hours, rem = divmod(rem, 3600)
    minutes, seconds = divmod(rem, 60)
    txt = '{:02d}:{:02d}:{:02d}'.format(hours, minutes, seconds)
    return '{}d'.format(days) + txt if days > 0 else txt


def print_log(txt, end='\n'):
    print(datetime2str() + ': ' + txt, end=end)
    sys.stdout.flush()


def print_tty(txt, end=''):
    if sys.stdout.isatty():
        print(txt, end=end)
        sys.stdout.flush()


def is_int(v):
    """True for text that reads as a whole number (percivaltts.py:131-135) -- including, as in the reference, an integral decimal
    such as '1.0' or '100.': signs are dropped from the left, then zeros and one dot from the right, and what remains must be
    digits only; a text holding '..' is tested as it stands."""
    v = str(v).strip()
    if v == '0':
        return True
    if '..' not in v:
        v = v.lstrip('-+').rstrip('0').rstrip('.')
    return v.isdigit()
