from . import utils  # noqa: F401
def main(*a, **k):
    def deco(f):
        return f
    return deco
