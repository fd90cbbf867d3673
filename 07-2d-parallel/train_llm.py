#!/usr/bin/env python
"""Chapter 07: 2D parallel, FSDP over dp x tensor parallel over tp (--tp).

Thin entry point: flags, loop and on-disk layout come from the shared driver
(``b200train/cli/train_llm.py``) with this chapter's strategy and defaults pinned.
Reference script: ``07-2d-parallel/train_llm.py``.
"""
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")))

from b200train.cli.train_llm import main  # noqa: E402

if __name__ == "__main__":
    main("07-2d-parallel")
