#!/usr/bin/env python3
"""Same script name and flags as the reference's ``02_development/compilation_optimization.py`` (--base_dir, --dtype
{fp32,bf16}, --repeat); see :mod:`hyperion_b200.bench.fusion`.  Unlike the reference it has a ``main()`` (nothing runs
at import) and sets no ROCm/MIOpen environment variables."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent))

from hyperion_b200.bench.fusion import main

if __name__ == "__main__":
    sys.exit(main())
