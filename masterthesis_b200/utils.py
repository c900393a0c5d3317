"""Host helpers mirrored from the reference's src/utils.py (AttributeDict semantics matter for
the Model API: nets/optimizers/losses are reachable both as keys and as attributes)."""
from collections import OrderedDict
from inspect import isclass


class AttributeDict(OrderedDict):
    """OrderedDict whose items are also attributes; a missing attribute reads as None
    (reference utils.py:122-148)."""

    def __init__(self, *args, **kwargs):
        super().__init__()
        for a in args:
            if isinstance(a, dict):
                for k, v in a.items():
                    self[k] = v
        for k, v in kwargs.items():
            self[k] = v

    def __getattr__(self, name):
        if name.startswith('_OrderedDict__') or name.startswith('__'):
            raise AttributeError(name)
        return self.get(name)

    def __setattr__(self, key, value):
        if key.startswith('_OrderedDict__'):
            return super().__setattr__(key, value)
        self[key] = value

    def __delattr__(self, item):
        del self[item]


def module_to_dict(module, exclude=()):
    """{class name: class} of a module (reference utils.py:29-32) — how `--model` is resolved."""
    return {n: getattr(module, n) for n in dir(module)
            if n not in exclude and isclass(getattr(module, n)) and getattr(module, n) not in exclude}
