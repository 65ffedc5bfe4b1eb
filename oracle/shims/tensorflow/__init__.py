"""Empty stand-in: helper.py:8 imports tensorflow at module level but only set_seed
(helper.py:36-41) touches it, and the oracle never calls set_seed."""
