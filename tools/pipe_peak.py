#!/usr/bin/env python3
"""Prints the integer-pipe microbenchmark (tools/pipe_peak.cu) as JSON: T lane-instr/s."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from bench import measure_pipe_peak  # noqa: E402

if __name__ == "__main__":
    print(json.dumps(measure_pipe_peak(0), indent=1))
