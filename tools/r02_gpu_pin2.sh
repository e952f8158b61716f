#!/bin/bash
timeout 600 python -m pytest tests/test_trilinear_pin.py -m gpu -q 2>&1 | tail -4
