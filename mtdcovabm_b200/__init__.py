"""mtdcovabm_b200: B200 (sm_100a) implementation of mtdcovabm's per-day contact network -> transmission -> SEIR state step."""
import os as _os

# A rank of a sharded run drives about a dozen CUDA streams (day pipeline, size-class streams of K1, state chain, copies).
# The driver maps streams onto 8 hardware queues by default, which makes independent streams wait for each other; with few
# cells per rank that false ordering is on the critical path (8 GPUs: 1.04 -> 0.93 ms/day, profiles/r02_k1_schedule_ab.txt).
# Only effective if set before the process creates its CUDA context; a value set by the user wins.
if int(_os.environ.get("WORLD_SIZE", "1")) > 1:
    _os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
