"""Progress prints of the reference scripts (masks_to_be_merged, detected_ratio_thres_value, ...).
The drop-in CLIs turn them on so stdout reads like the reference's; library callers get silence
(formatting tensors for a print costs device syncs and milliseconds)."""
VERBOSE = False


def log(*args):
    if VERBOSE:
        print(*args)
