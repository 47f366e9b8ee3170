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
