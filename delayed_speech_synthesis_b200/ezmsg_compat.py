"""The slice of ezmsg 3.0 that `HighGammaActivity` needs (Unit/Settings/State, stream descriptors, the
subscriber/publisher decorators) and `TimeSeriesMessage`.  Real ezmsg is used when importable (the reference imports
`ezmsg.core as ez` and `ezmsg.eeg.eegmessage.TimeSeriesMessage`, local/units.py:13,15); it is not installed in this
image and not vendored by the reference, so a minimal stand-in keeps the unit importable and testable."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Optional

try:  # pragma: no cover - exercised only where ezmsg is installed
    import ezmsg.core as ez  # type: ignore
    from ezmsg.eeg.eegmessage import TimeSeriesMessage  # type: ignore

    HAVE_EZMSG = True
except ImportError:
    HAVE_EZMSG = False

    class _Namespace:
        pass

    ez = _Namespace()

    class _SettingsMeta(type):
        """Turns `class X(ez.Settings): a: int; b: float = 1.0` into a frozen dataclass, like ezmsg does."""

        def __new__(mcls, name, bases, ns, **kw):
            cls = super().__new__(mcls, name, bases, ns, **kw)
            if bases:
                cls = dataclass(frozen=True)(cls)
            return cls

    class Settings(metaclass=_SettingsMeta):
        pass

    class _StateMeta(type):
        def __new__(mcls, name, bases, ns, **kw):
            cls = super().__new__(mcls, name, bases, ns, **kw)
            if bases:
                cls = dataclass(cls)
            return cls

    class State(metaclass=_StateMeta):
        pass

    class _Stream:
        def __init__(self, msg_type: Any = Any):
            self.msg_type = msg_type
            self.name: Optional[str] = None

        def __set_name__(self, owner, name):
            self.name = name

    class InputStream(_Stream):
        pass

    class OutputStream(_Stream):
        pass

    def subscriber(stream):
        def deco(fn):
            fn.__ez_subscribes__ = stream
            return fn
        return deco

    def publisher(stream):
        def deco(fn):
            fn.__ez_publishes__ = getattr(fn, "__ez_publishes__", []) + [stream]
            return fn
        return deco

    class Unit:
        """SETTINGS/STATE are instantiated from the class annotations; `apply_settings` then `initialize`, as
        ezmsg's `Unit.setup()` does before the graph starts."""

        def __init__(self, settings=None):
            import typing

            try:
                ann = typing.get_type_hints(type(self))
            except Exception:
                ann = {}
            state_t = ann.get("STATE")
            self.STATE = state_t() if isinstance(state_t, type) else None
            self.SETTINGS = None
            if settings is not None:
                self.apply_settings(settings)

        def apply_settings(self, settings) -> None:
            self.SETTINGS = settings

        def initialize(self) -> None:
            pass

        def setup(self) -> None:
            self.initialize()

    for _n, _o in dict(Settings=Settings, State=State, Unit=Unit, InputStream=InputStream, OutputStream=OutputStream,
                       subscriber=subscriber, publisher=publisher).items():
        setattr(ez, _n, _o)

    @dataclass
    class TimeSeriesMessage:
        """ezmsg.eeg's message: a (time, channel) array, its sampling rate and which axis is time."""
        data: Any
        fs: float = 1.0
        time_dim: int = 0
