"""tests/ref_mpl — TEST INFRASTRUCTURE: a headless stand-in for matplotlib (absent from this image, and there is no graphics stack), so
that the plotting lines of the reference's python/examples scripts run as no-ops when tools/run_ref_python_examples.sh executes those
scripts unmodified against this repository's `goofit` package. Every attribute exists, every call returns another stand-in."""
import sys


class _Anything:
    def __getattr__(self, name):
        return _Anything()

    def __call__(self, *args, **kwargs):
        return _Anything()

    def __iter__(self):
        return iter([_Anything(), _Anything()])

    def __getitem__(self, item):
        return _Anything()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


def use(*args, **kwargs):
    pass


pyplot = _Anything()
sys.modules[__name__ + ".pyplot"] = pyplot
