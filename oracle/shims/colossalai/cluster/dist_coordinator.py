class DistCoordinator:  # empty stand-in (import only, utils/misc.py:11)
    pass
