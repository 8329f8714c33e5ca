class DeprecatedException(Exception):
    pass
