"""Stand-in for `jsonpickle` (missing in this image), covering exactly what the reference runner uses:
`encode(Metadata, indent=2)` / `decode(str)` in ProgressManager/Output/JSONOutputManager.py:9-16, producing the
shape of the shipped experiment/experiments_output/new_runner_experiment/metadata.json:1-5."""
import base64
import importlib
import json


def _enc(v):
    if isinstance(v, (bytes, bytearray)):
        return {"py/b64": base64.b64encode(bytes(v)).decode()}
    if isinstance(v, (str, int, float, bool)) or v is None:
        return v
    if isinstance(v, (list, tuple)):
        return [_enc(x) for x in v]
    if isinstance(v, dict):
        return {k: _enc(x) for k, x in v.items()}
    d = {"py/object": f"{type(v).__module__}.{type(v).__qualname__}"}
    d.update({k: _enc(x) for k, x in vars(v).items()})
    return d


def _dec(v):
    if isinstance(v, list):
        return [_dec(x) for x in v]
    if isinstance(v, dict):
        if set(v) == {"py/b64"}:
            return base64.b64decode(v["py/b64"])
        if "py/object" in v:
            mod, _, name = v["py/object"].rpartition(".")
            cls = getattr(importlib.import_module(mod), name)
            obj = cls.__new__(cls)
            for k, x in v.items():
                if k != "py/object":
                    setattr(obj, k, _dec(x))
            return obj
        return {k: _dec(x) for k, x in v.items()}
    return v


def encode(obj, indent=None, **_):
    return json.dumps(_enc(obj), indent=indent)


def decode(s, **_):
    return _dec(json.loads(s))
