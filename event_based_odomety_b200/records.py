"""numpy mirrors of the plain-C records in include/ebt_types.h.

The dtypes are built with ``align=True`` so their layout equals the C structs;
``tests/test_abi.py`` checks every size against ``ebt_sizeof_*`` exported by
the library.
"""
import numpy as np

PATCH_EXTENT = 12
PATCH_SIDE = 25
PATCH_PIXELS = 625

# common::EventSample -> 16-byte wire record (include/ebt_types.h: ebt_event)
EVENT = np.dtype(
    [("t_us", "<i8"), ("x", "<u2"), ("y", "<u2"), ("polarity", "i1"), ("pad_", "i1", (3,))],
    align=True,
)

TRACK = np.dtype(
    [
        ("track_id", "<i4"), ("lost", "<i4"), ("init", "<i4"), ("num_of_events", "<i4"),
        ("counter", "<i4"), ("window_size", "<i4"), ("n_optimizations", "<i4"),
        ("n_trajectory", "<i4"),
        ("rect_x", "<f8"), ("rect_y", "<f8"), ("init_x", "<f8"), ("init_y", "<f8"),
        ("warp", "<f8", (4,)), ("flow_dir", "<f8"), ("last_cost", "<f8"),
        ("current_ts", "<i8"), ("time_last_update", "<i8"), ("time_without_update", "<i8"),
        ("init_time", "<i8"),
    ],
    align=True,
)

OPT_RECORD = np.dtype(
    [
        ("track_id", "<i4"), ("n_successful", "<i4"), ("n_unsuccessful", "<i4"),
        ("termination", "<i4"), ("lost", "<i4"), ("num_of_events", "<i4"),
        ("ts", "<i8"), ("event_index", "<i8"),
        ("x", "<f8"), ("y", "<f8"), ("final_cost", "<f8"),
        ("warp", "<f8", (4,)), ("flow_dir", "<f8"),
    ],
    align=True,
)

PROBLEM = np.dtype(
    [
        ("init_x", "<f8"), ("init_y", "<f8"), ("warp", "<f8", (4,)), ("flow_dir", "<f8"),
        ("n_events", "<i4"), ("event_offset", "<i4"), ("frame_slot", "<i4"), ("pad_", "<i4"),
    ],
    align=True,
)

SOLUTION = np.dtype(
    [
        ("warp", "<f8", (4,)), ("flow_dir", "<f8"), ("final_cost", "<f8"),
        ("x", "<f8"), ("y", "<f8"),
        ("n_successful", "<i4"), ("n_unsuccessful", "<i4"), ("termination", "<i4"),
        ("n_passes", "<i4"), ("ts", "<i8"),
    ],
    align=True,
)

assert EVENT.itemsize == 16
assert PROBLEM.itemsize == 72
