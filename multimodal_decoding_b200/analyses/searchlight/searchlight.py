"""Alias for the path the reference's README and BASELINE.json name (`analyses/searchlight/searchlight.py`,
README.md:18); the module lives at `analyses/decoding/searchlight/searchlight.py` like the reference's real file."""
from ..decoding.searchlight.searchlight import *          # noqa: F401,F403
from ..decoding.searchlight.searchlight import get_args, run, SEARCHLIGHT_OUT_DIR   # noqa: F401

if __name__ == "__main__":
    import os
    os.makedirs(SEARCHLIGHT_OUT_DIR, exist_ok=True)
    run(get_args())
