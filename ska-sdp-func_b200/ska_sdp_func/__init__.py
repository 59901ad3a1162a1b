"""ska_sdp_func: Python bindings of the B200-native drop-in for the
sdp_GridderUvwEsFft path (same names as the reference package's
``ska_sdp_func.grid_data.GridderUvwEsFft`` and ``ska_sdp_func.utility``)."""
