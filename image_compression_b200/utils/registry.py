"""Name -> class registry with the interface the reference uses (utils/registry.py:23-92):
`R = Registry("X")`, `@R.register()` or `R.register(obj)`, `R.get(name)`, `name in R`."""


class Registry(dict):
    def __init__(self, name):
        super().__init__()
        self._name = name

    def register(self, obj=None):
        if obj is None:          # decorator form
            return lambda o: (self.register(o), o)[1]
        key = obj.__name__
        assert key not in self, f"An object named '{key}' was already registered in '{self._name}' registry!"
        self[key] = obj
        return None

    def get(self, name):
        try:
            return self[name]
        except KeyError:
            raise KeyError(f"No object named '{name}' found in '{self._name}' registry!") from None
