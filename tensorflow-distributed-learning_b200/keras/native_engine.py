"""Placeholder replaced below by the native step-program engine."""


def try_build(model):
    return None
