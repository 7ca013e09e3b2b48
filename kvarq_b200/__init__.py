"""kvarq_b200 -- B200-native (sm_100a) read-scanning engine behind KvarQ's ``kvarq.engine`` API.

``from kvarq_b200 import engine`` is the drop-in for ``from kvarq import engine``.
"""
VERSION = '0.1.0'
