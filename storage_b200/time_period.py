"""Minimal time-period value types: the subset of Cmdty.TimePeriodValueTypes (un-vendored NuGet dependency of the
reference) that the LSMC path touches — Offset / OffsetFrom / EnumerateTo / First<Day>() / comparison
(call sites: StorageHelper.cs:45-60, LsmcStorageValuation.cs:92,134,157, LsmcValuationParameters.cs:177-178).
"""
import datetime as _dt
import functools


@functools.total_ordering
class Day:
    """A calendar day; the granularity every reference test and sample of the LSMC path uses."""
    __slots__ = ("ordinal",)

    def __init__(self, year, month=None, day=None):
        if month is None:
            if isinstance(year, Day):
                self.ordinal = year.ordinal
            elif isinstance(year, (_dt.date, _dt.datetime)):
                self.ordinal = _dt.date(year.year, year.month, year.day).toordinal()
            else:
                self.ordinal = int(year)
        else:
            self.ordinal = _dt.date(year, month, day).toordinal()

    @property
    def Start(self):
        return _dt.date.fromordinal(self.ordinal)

    def Offset(self, n):
        return Day(self.ordinal + int(n))

    def OffsetFrom(self, other):
        return self.ordinal - other.ordinal

    def First(self):
        return self

    def EnumerateTo(self, end):
        return [Day(o) for o in range(self.ordinal, end.ordinal + 1)]

    def __add__(self, n):
        return self.Offset(n)

    def __sub__(self, other):
        if isinstance(other, Day):
            return self.ordinal - other.ordinal
        return self.Offset(-int(other))

    def __eq__(self, other):
        return isinstance(other, Day) and self.ordinal == other.ordinal

    def __lt__(self, other):
        return self.ordinal < other.ordinal

    def __hash__(self):
        return hash(self.ordinal)

    def __repr__(self):
        return f"Day({self.Start.isoformat()})"


@functools.total_ordering
class Month:
    __slots__ = ("index",)

    def __init__(self, year, month=None):
        self.index = int(year) if month is None else year * 12 + (month - 1)

    @staticmethod
    def FromDateTime(d):
        return Month(d.year, d.month)

    def Offset(self, n):
        return Month(self.index + int(n))

    def First(self):
        return Day(self.index // 12, self.index % 12 + 1, 1)

    def __eq__(self, other):
        return isinstance(other, Month) and self.index == other.index

    def __lt__(self, other):
        return self.index < other.index

    def __hash__(self):
        return hash(("M", self.index))

    def __repr__(self):
        return f"Month({self.index // 12}-{self.index % 12 + 1:02d})"


def settle_on_20th_of_following_month(delivery_day):
    """The settlement rule of the reference's fixtures and samples:
    Month.FromDateTime(d.Start).Offset(1).First<Day>() + 19 (tests/Cmdty.Storage.Test/Lsmc/LsmcStorageValuationTest.cs:154)."""
    return Month.FromDateTime(delivery_day.Start).Offset(1).First() + 19
