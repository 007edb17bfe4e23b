"""The two tunables of the hot path, with the reference's defaults (constants.py:10-11).
They are read at call time, so assigning to them (as one does with the reference's
module) changes the ROI size / light-grid step of the next call."""
ROI_DIAMETER = 400
INTERPOLATION_PARAM = 0.02
