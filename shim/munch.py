"""Minimal stand-in for `munch.Munch`, which the reference's src/utils.py:1,8-13 imports to read ./src/config.yaml and which
is not installed in this image.  Only what the reference uses: attribute access on nested dicts and the
`fromYAML` constructor.  Put this directory on sys.path ahead of site-packages."""


class Munch(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    def __delattr__(self, k):
        try:
            del self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    @classmethod
    def _wrap(cls, x):
        if isinstance(x, dict):
            return cls({k: cls._wrap(v) for k, v in x.items()})
        if isinstance(x, list):
            return [cls._wrap(v) for v in x]
        return x

    @classmethod
    def fromDict(cls, d):
        return cls._wrap(d)

    @classmethod
    def fromYAML(cls, stream, *a, **k):
        import yaml
        return cls._wrap(yaml.safe_load(stream))

    def toDict(self):
        return {k: (v.toDict() if isinstance(v, Munch) else v) for k, v in self.items()}


munchify = Munch.fromDict
