"""Stand-in for the absent `func_timeout` package (test infrastructure only).
Used as a decorator at /root/reference/UTIL/config_args.py:267."""


def func_set_timeout(_t):
    return lambda f: f


class FunctionTimedOut(Exception):
    pass


class exceptions(object):
    FunctionTimedOut = FunctionTimedOut
