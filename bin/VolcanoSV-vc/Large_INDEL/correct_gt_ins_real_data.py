#!/usr/bin/env python3
"""Drop-in for the reference's bin/VolcanoSV-vc/Large_INDEL/correct_gt_ins_real_data.py: same name, same argv, same files; the arithmetic runs in
libvolcano_b200.so (CUDA, sm_100a) through volcanosv_b200.cli."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.realpath(__file__)), "..", "..", ".."))
from volcanosv_b200 import cli  # noqa: E402

cli.main(['correct_gt_ins_real_data.py'] + sys.argv[1:])
