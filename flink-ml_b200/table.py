"""A minimal stand-in for Flink's Table on this side of the boundary: named columns of equal
length.  A vector column is an n x d FP64 array (or a sequence of vectors); other columns are
carried through transform() untouched, as Row.join does (KMeansModel.java:143)."""
import numpy as np


class Table:
    def __init__(self, columns):
        self._cols = dict(columns)
        n = None
        for name, col in self._cols.items():
            m = len(col)
            if n is None:
                n = m
            elif m != n:
                raise ValueError("column %r has %d rows, expected %d" % (name, m, n))
        self._n = n or 0

    @staticmethod
    def from_rows(rows, name="features"):
        return Table({name: rows})

    def get_column_names(self):
        return list(self._cols.keys())

    def column(self, name):
        if name not in self._cols:
            raise KeyError("no column %r; columns are %s" % (name, self.get_column_names()))
        return self._cols[name]

    def num_rows(self):
        return self._n

    def with_column(self, name, values):
        cols = dict(self._cols)
        cols[name] = values
        return Table(cols)

    def as_(self, *names):
        if len(names) != len(self._cols):
            raise ValueError("expected %d names" % len(self._cols))
        return Table(dict(zip(names, self._cols.values())))

    def __len__(self):
        return self._n

    def __repr__(self):
        return "Table(%s, rows=%d)" % (self.get_column_names(), self._n)


def is_vector_column(col):
    return isinstance(col, np.ndarray) and col.ndim == 2
