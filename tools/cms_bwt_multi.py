#!/usr/bin/env python3
"""Multi-GPU cms_bwt: torchrun --nproc-per-node N tools/cms_bwt_multi.py [-p P] [-r] [-o out] list.txt"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cms_bwt_b200 import multi
multi.main()
