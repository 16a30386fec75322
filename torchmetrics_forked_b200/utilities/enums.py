"""Task enums used by the `task=` wrapper classes (reference: utilities/enums.py:19-53, 108-156)."""
from enum import Enum


class _TaskEnum(str, Enum):
    """Compares equal to its lower-case string value; ``from_str`` is a case-insensitive lookup by member name."""

    @classmethod
    def from_str(cls, value: str) -> "_TaskEnum":
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


class ClassificationTask(_TaskEnum):
    """The three classification tasks."""

    BINARY = "binary"
    MULTICLASS = "multiclass"
    MULTILABEL = "multilabel"


class ClassificationTaskNoMultilabel(_TaskEnum):
    """Tasks of the metrics that have no multilabel form (Cohen's kappa)."""

    BINARY = "binary"
    MULTICLASS = "multiclass"
