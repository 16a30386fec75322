"""Task enum used by the `task=` wrapper classes (reference: utilities/enums.py:19-53, 108-122)."""
from enum import Enum


class ClassificationTask(str, Enum):
    """The three classification tasks; compares equal to its lower-case string value."""

    BINARY = "binary"
    MULTICLASS = "multiclass"
    MULTILABEL = "multilabel"

    @classmethod
    def from_str(cls, value: str) -> "ClassificationTask":
        """Case-insensitive lookup by member name; dashes count as underscores."""
        if isinstance(value, cls):
            return value
        key = str(value).replace("-", "_").lower()
        for member in cls:
            if member.name.lower() == key:
                return member
        allowed = [m.name.lower() for m in cls]
        raise ValueError(f"Invalid Classification: expected one of {allowed}, but got {value}.")

    def __eq__(self, other: object) -> bool:
        if isinstance(other, Enum):
            other = other.value
        return self.value.lower() == str(other).lower()

    def __hash__(self) -> int:
        return hash(self.value.lower())

    def __str__(self) -> str:
        return self.value
