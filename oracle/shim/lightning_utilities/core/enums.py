"""`StrEnum` with case-insensitive lookup and comparison (stand-in, see package docstring)."""
from enum import Enum
from typing import List, Optional


class StrEnum(str, Enum):
    @classmethod
    def from_str(cls, value: str, source: str = "key") -> "StrEnum":
        if source in ("key", "any"):
            for name, member in cls.__members__.items():
                if name.lower() == value.lower():
                    return member
        if source in ("value", "any"):
            for member in cls:
                if member.value.lower() == value.lower():
                    return member
        raise ValueError(f"Invalid match: expected one of {cls._allowed_matches(source)}, but got {value}.")

    @classmethod
    def try_from_str(cls, value: str, source: str = "key") -> Optional["StrEnum"]:
        try:
            return cls.from_str(value, source)
        except ValueError:
            return None

    @classmethod
    def _allowed_matches(cls, source: str) -> List[str]:
        keys = [m.lower() for m in cls.__members__]
        vals = [m.value.lower() for m in cls]
        if source == "key":
            return keys
        if source == "value":
            return vals
        return keys + vals

    def __eq__(self, other: object) -> bool:
        if isinstance(other, Enum):
            other = other.value
        return self.value.lower() == str(other).lower()

    def __hash__(self) -> int:
        return hash(self.value.lower())
