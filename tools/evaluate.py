#!/usr/bin/env python
"""Headless counterpart of the reference's `evaluate <config.json> <output_dir>` (src/evaluate.cu:118-310).

    python tools/evaluate.py configs/synthetic_ico.json out/ [--width W --height H]
"""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

if __name__ == "__main__":
    sys.exit(importlib.import_module("neural-renderer_b200.evaluate").main())
