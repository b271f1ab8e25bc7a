#!/usr/bin/env python
"""Drop-in for `yolov5/detect.py` [REF README.md:77] running on the B200-native engine.
See aquaculture_b200/detect_cli.py."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from aquaculture_b200.detect_cli import main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main())
