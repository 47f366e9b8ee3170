"""Process-wide single instances (dlrover/python/common/singleton.py:31-48)."""
import threading


class Singleton(object):
    _instance = None
    _guard = threading.Lock()

    @classmethod
    def singleton_instance(cls, *args, **kwargs):
        if cls.__dict__.get("_instance") is None:
            with Singleton._guard:
                if cls.__dict__.get("_instance") is None:
                    cls._instance = cls(*args, **kwargs)
        return cls._instance


def singleton(cls):
    """Class decorator: calling the class returns the one instance built by
    the first call (dlrover/python/common/singleton.py:18-28; ATorch decorates
    its checkpoint engine with it)."""
    guard = threading.Lock()
    box = []

    def get(*args, **kwargs):
        with guard:
            if not box:
                box.append(cls(*args, **kwargs))
        return box[0]

    get.__wrapped__ = cls
    return get
