#!/usr/bin/env python
"""Same place as the reference's logs/plot.py; the implementation is tools/plot.py (headless, writes a file)."""
import os
import runpy
import sys

if __name__ == '__main__':
    here = os.path.dirname(os.path.abspath(__file__))
    sys.argv[0] = os.path.join(here, '..', 'tools', 'plot.py')
    runpy.run_path(sys.argv[0], run_name='__main__')
