"""`apply_to_collection`: map a function over the leaves of a nested collection (stand-in, see package docstring)."""
from collections import OrderedDict, defaultdict
from typing import Any, Callable, Tuple, Union


def apply_to_collection(data: Any, dtype: Union[type, Tuple[type, ...]], function: Callable, *args: Any, **kwargs: Any) -> Any:
    if isinstance(data, dtype):
        return function(data, *args, **kwargs)
    if isinstance(data, defaultdict):
        return defaultdict(data.default_factory, {k: apply_to_collection(v, dtype, function, *args, **kwargs) for k, v in data.items()})
    if isinstance(data, (dict, OrderedDict)):
        return type(data)((k, apply_to_collection(v, dtype, function, *args, **kwargs)) for k, v in data.items())
    if isinstance(data, tuple) and hasattr(data, "_fields"):
        return type(data)(*(apply_to_collection(v, dtype, function, *args, **kwargs) for v in data))
    if isinstance(data, (list, tuple)):
        return type(data)(apply_to_collection(v, dtype, function, *args, **kwargs) for v in data)
    return data
