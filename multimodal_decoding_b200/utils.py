"""Constants of the reference's `utils.py` that the hot path and its CLIs use (utils.py:10-50).  The nibabel /
matplotlib helpers of that file (GIfTI export, image cropping) are out of scope."""
import os

DATA_DIR = os.path.expanduser(os.environ.get("MMD_DATA_DIR", "~/data/multimodal_decoding"))   # utils.py:12
RESULTS_FILE = "results.p"                                                                     # utils.py:19
FMRI_DATA_DIR = os.path.join(DATA_DIR, "fmri")
FMRI_BIDS_DATA_DIR = os.path.join(FMRI_DATA_DIR, "bids")
FMRI_BETAS_DIR = os.path.join(FMRI_BIDS_DATA_DIR, "derivatives", "betas")                      # utils.py:31
FMRI_BETAS_SURFACE_DIR = os.path.join(FMRI_BETAS_DIR, "surface")                               # utils.py:33
RIDGE_DECODER_OUT_DIR = os.path.join(DATA_DIR, "whole_brain_decoding/")                        # utils.py:41

SUBJECTS = ['sub-01', 'sub-02', 'sub-03', 'sub-04', 'sub-05', 'sub-07']                        # utils.py:43
HEMIS = ['left', 'right']
DEFAULT_MODEL = "imagebind"                                                                    # utils.py:49
DEFAULT_RESOLUTION = "fsaverage7"                                                              # utils.py:50
