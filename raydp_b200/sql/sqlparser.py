"""A small SQL front end: ``spark.sql(...)``, ``df.filter("...")``, ``df.selectExpr(...)``, ``F.expr``.

The reference leaves SQL to Spark's JVM parser (``doc/spark_on_ray.md:187-202`` uses
``spark.sql("select * from db.xxx")`` with ``enable_hive=True``).  There is no JVM here, so the
subset people actually type is parsed directly into the DataFrame API:

    SELECT [DISTINCT] <exprs | *> FROM <view | db.table>
        [WHERE <predicate>] [GROUP BY <cols>] [HAVING <predicate>]
        [ORDER BY <expr> [ASC|DESC], ...] [LIMIT n]
    SHOW TABLES [IN db] | SHOW DATABASES | CREATE DATABASE [IF NOT EXISTS] db
    DROP TABLE [IF EXISTS] t | DROP VIEW [IF EXISTS] v
    CREATE TABLE t [USING parquet] AS SELECT ... | INSERT INTO t SELECT ...

Expressions: literals, column names, ``+ - * / %``, comparisons, ``AND / OR / NOT``, ``IS [NOT] NULL``,
``IN (...)``, ``BETWEEN a AND b``, ``CAST(x AS type)``, ``CASE WHEN ... THEN ... [ELSE ...] END`` and calls
of anything in :mod:`raydp_b200.sql.functions` (aggregates included).  ``FROM a [INNER|LEFT|RIGHT|FULL|
CROSS|LEFT SEMI|LEFT ANTI] JOIN b ON a.k = b.k [AND ...] | USING (k)`` is an equi-join (column qualifiers are
dropped: a join key named differently on the two sides keeps the left name).  Sub-queries and window
clauses are not parsed; use the DataFrame API for those.
"""
from __future__ import annotations

import re
from typing import Any, Dict, List, Optional, Tuple

from raydp_b200.sql import functions as F
from raydp_b200.sql.column import Column

_TOKEN = re.compile(r"""
    \s*(?:
      (?P<num>\d+\.\d*(?:[eE][+-]?\d+)?|\.\d+|\d+(?:[eE][+-]?\d+)?)
    | (?P<str>'(?:[^']|'')*'|"(?:[^"]|"")*")
    | (?P<bq>`[^`]+`)
    | (?P<id>[A-Za-z_][A-Za-z_0-9]*)
    | (?P<op><=>|<=|>=|<>|!=|==|\|\||[-+*/%(),.<>=\[\]])
    )""", re.X)

_AGGREGATES = {"count", "sum", "avg", "mean", "min", "max", "first", "stddev", "variance", "countdistinct"}
_KEYWORDS = {"select", "from", "where", "group", "by", "having", "order", "limit", "as", "and", "or", "not",
             "is", "null", "in", "between", "case", "when", "then", "else", "end", "cast", "distinct", "asc",
             "desc", "true", "false", "like", "join", "inner", "left", "right", "full", "cross", "on", "using",
             "union", "intersect", "except", "minus", "over", "rlike", "ilike", "regexp", "exists", "with",
             "partition", "rows", "range", "window", "outer", "semi", "anti", "natural", "lateral"}


class SqlError(ValueError):
    pass


def tokenize(text: str) -> List[Tuple[str, str]]:
    out, pos = [], 0
    text = re.sub(r"--[^\n]*", " ", text)            # line comments
    text = text.strip().rstrip(";")
    while pos < len(text):
        m = _TOKEN.match(text, pos)
        if m is None or m.end() == pos:
            if text[pos:].strip() == "":
                break
            raise SqlError(f"cannot tokenize SQL at: {text[pos:pos + 20]!r}")
        pos = m.end()
        kind = m.lastgroup
        value = m.group(kind)
        if kind == "str":
            q = value[0]
            value = value[1:-1].replace(q + q, q)
        elif kind == "bq":
            kind, value = "qid", value[1:-1]
        out.append((kind, value))
    return out


class _Parser:
    def __init__(self, tokens: List[Tuple[str, str]], session=None, ctes: Optional[Dict[str, Any]] = None):
        self.t = tokens
        self.i = 0
        self.session = session
        self.ctes: Dict[str, Any] = dict(ctes or {})

    def _subquery(self):
        """``( SELECT ... )`` with the opening parenthesis already consumed; consumes the closing one."""
        if self.session is None:
            raise SqlError("sub-queries need a session: use spark.sql(...)")
        try:
            df = _parse_query(self, self.session)
            self.expect_op(")")
            df.columns                      # resolves the plan: an outer reference cannot be found here
        except KeyError as exc:
            raise SqlError(f"correlated sub-queries are not supported (the sub-query refers to a column of the outer "
                           f"query: {exc}); rewrite it as a join (LEFT SEMI / LEFT ANTI for EXISTS / NOT EXISTS)") from exc
        return df

    # ---- token helpers
    def peek(self, k: int = 0) -> Tuple[str, str]:
        j = self.i + k
        return self.t[j] if j < len(self.t) else ("eof", "")

    def kw(self, *words: str) -> bool:
        """True (and consume) if the next tokens are the given keywords."""
        for k, w in enumerate(words):
            kind, v = self.peek(k)
            if kind != "id" or v.lower() != w:
                return False
        self.i += len(words)
        return True

    def at_kw(self, word: str) -> bool:
        kind, v = self.peek()
        return kind == "id" and v.lower() == word

    def op(self, sym: str) -> bool:
        if self.peek() == ("op", sym):
            self.i += 1
            return True
        return False

    def expect_op(self, sym: str) -> None:
        if not self.op(sym):
            raise SqlError(f"expected {sym!r}, found {self.peek()[1]!r}")

    def name(self) -> str:
        kind, v = self.peek()
        if kind not in ("id", "qid"):
            raise SqlError(f"expected a name, found {v!r}")
        self.i += 1
        return v

    def done(self) -> bool:
        return self.i >= len(self.t)

    # ---- expressions (precedence: OR < AND < NOT < comparison < additive < multiplicative < unary)
    def expr(self) -> Column:
        left = self.and_expr()
        while self.kw("or"):
            left = left | self.and_expr()
        return left

    def and_expr(self) -> Column:
        left = self.not_expr()
        while self.kw("and"):
            left = left & self.not_expr()
        return left

    def not_expr(self) -> Column:
        if self.kw("not"):
            return ~self.not_expr()
        return self.comparison()

    def comparison(self) -> Column:
        left = self.additive()
        while True:
            if self.kw("is", "not", "null"):
                left = left.isNotNull()
            elif self.kw("is", "null"):
                left = left.isNull()
            elif self.kw("is", "not", "distinct", "from"):
                left = left.eqNullSafe(self.additive())
            elif self.kw("is", "distinct", "from"):
                left = ~left.eqNullSafe(self.additive())
            elif self.kw("is", "not", "true") or self.kw("is", "false"):
                left = F.coalesce(left, F.lit(False)) == F.lit(False)
            elif self.kw("is", "true") or self.kw("is", "not", "false"):
                left = F.coalesce(left, F.lit(False)) == F.lit(True)
            elif self.at_kw("not") and self.peek(1)[1].lower() in ("in", "between"):
                self.i += 1
                left = ~self._in_or_between(left)
            elif self.at_kw("in") or self.at_kw("between"):
                left = self._in_or_between(left)
            elif self.at_kw("not") and self.peek(1)[1].lower() in ("like", "rlike", "ilike", "regexp"):
                self.i += 1
                left = ~self._like(left)
            elif self.at_kw("like") or self.at_kw("rlike") or self.at_kw("ilike") or self.at_kw("regexp"):
                left = self._like(left)
            elif self.op("<=>"):
                left = left.eqNullSafe(self.additive())
            else:
                kind, v = self.peek()
                if kind == "op" and v in ("=", "==", "<>", "!=", "<", "<=", ">", ">="):
                    self.i += 1
                    right = self.additive()
                    left = {"=": left == right, "==": left == right, "<>": left != right, "!=": left != right,
                            "<": left < right, "<=": left <= right, ">": left > right, ">=": left >= right}[v]
                else:
                    return left

    def _like(self, left: Column) -> Column:
        word = self.name().lower()
        pattern = self.additive()
        if pattern._kind != "lit":
            raise SqlError(f"{word.upper()} needs a literal pattern")
        text = pattern._args[0]
        if self.kw("escape"):
            self.i += 1
        return {"like": left.like, "ilike": left.ilike, "rlike": left.rlike, "regexp": left.rlike}[word](text)

    def _in_or_between(self, left: Column) -> Column:
        if self.kw("in"):
            self.expect_op("(")
            if self.at_kw("select") or self.at_kw("with"):
                sub = self._subquery()
                return left.isin(*[r[0] for r in sub.collect() if r[0] is not None])
            values = [self._literal_value()]
            while self.op(","):
                values.append(self._literal_value())
            self.expect_op(")")
            return left.isin(*values)
        self.kw("between")
        lo = self.additive()
        if not self.kw("and"):
            raise SqlError("BETWEEN needs AND")
        return left.between(lo, self.additive())

    def _literal_value(self) -> Any:
        kind, v = self.peek()
        sign = 1
        if kind == "op" and v == "-":
            self.i += 1
            sign = -1
            kind, v = self.peek()
        self.i += 1
        if kind == "num":
            return sign * (float(v) if any(c in v for c in ".eE") else int(v))
        if kind == "str":
            return v
        raise SqlError(f"expected a literal in the IN list, found {v!r}")

    def additive(self) -> Column:
        left = self.multiplicative()
        while True:
            if self.op("+"):
                right = self.multiplicative()
                if right._kind == "interval":
                    left = self._shift(left, right, 1)
                elif left._kind == "interval":
                    left = self._shift(right, left, 1)
                else:
                    left = left + right
            elif self.op("-"):
                right = self.multiplicative()
                left = self._shift(left, right, -1) if right._kind == "interval" else left - right
            elif self.op("||"):
                left = F.concat(left, self.multiplicative())
            else:
                return left

    def multiplicative(self) -> Column:
        left = self.unary()
        while True:
            if self.op("*"):
                left = left * self.unary()
            elif self.op("/"):
                left = left / self.unary()
            elif self.op("%"):
                left = left % self.unary()
            elif self.kw("div"):            # integral division: 7 div 2 = 3, -7 div 2 = -3 (truncates)
                left = (left / self.unary()).cast("long")
            else:
                return left

    def unary(self) -> Column:
        if self.op("-"):
            return -self.unary()
        if self.op("+"):
            return self.unary()
        return self.primary()

    def primary(self) -> Column:
        kind, v = self.peek()
        if kind == "num":
            self.i += 1
            return F.lit(float(v) if any(c in v for c in ".eE") else int(v))
        if kind == "str":
            self.i += 1
            return F.lit(v)
        if kind == "op" and v == "(":
            self.i += 1
            if self.at_kw("select") or self.at_kw("with"):
                sub = self._subquery()                       # scalar sub-query (uncorrelated): first cell
                rows = sub.limit(2).collect()
                if len(rows) > 1:
                    raise SqlError("a scalar sub-query returned more than one row")
                return F.lit(rows[0][0] if rows else None)
            inner = self.expr()
            self.expect_op(")")
            return self._postfix(inner)
        if kind == "qid":
            self.i += 1
            return F.col(v)
        if kind != "id":
            raise SqlError(f"unexpected {v!r} in expression")
        low = v.lower()
        if low in ("true", "false"):
            self.i += 1
            return F.lit(low == "true")
        if low == "null":
            self.i += 1
            return F.lit(None)
        if low == "case":
            return self._case()
        if low == "exists" and self.peek(1) == ("op", "("):
            self.i += 2
            return F.lit(self._subquery().limit(1).count() > 0)
        if low in ("date", "timestamp") and self.peek(1)[0] == "str":
            self.i += 2
            text = self.t[self.i - 1][1]
            return F.to_date(F.lit(text)) if low == "date" else F.to_timestamp(F.lit(text))
        if low == "interval":
            self.i += 1
            return self._interval()
        if low == "cast" and self.peek(1) == ("op", "("):
            self.i += 2
            inner = self.expr()
            if not self.kw("as"):
                raise SqlError("CAST needs AS <type>")
            type_name = self._type_text()
            self.expect_op(")")
            return self._postfix(inner.cast(type_name.lower()))
        if self.peek(1) == ("op", "("):
            return self._postfix(self._call())
        self.i += 1
        # qualified name: alias.column (resolved at evaluation: "<alias>.<column>" if the join kept it
        # apart, else the bare column) or struct_column.field
        while self.peek() == ("op", ".") and (self.peek(1)[0] in ("id", "qid") or self.peek(1) == ("op", "*")):
            self.i += 1
            if self.op("*"):
                return F.col(v + ".*")
            v = v + "." + self.name()
        return self._postfix(F.col(v))

    _INTERVAL_UNITS = {"year": (12, 0, 0), "month": (1, 0, 0), "week": (0, 7, 0), "day": (0, 1, 0),
                       "hour": (0, 0, 3600 * 10 ** 6), "minute": (0, 0, 60 * 10 ** 6), "second": (0, 0, 10 ** 6),
                       "millisecond": (0, 0, 1000), "microsecond": (0, 0, 1)}

    def _interval(self) -> Column:
        """``INTERVAL 3 DAY``, ``INTERVAL '3' DAY``, ``INTERVAL '1 day 2 hours'``, ``INTERVAL 1 YEAR 2 MONTHS``."""
        months = days = micros = 0

        def add(qty: float, unit: str) -> None:
            nonlocal months, days, micros
            key = unit.lower().rstrip("s")
            if key not in self._INTERVAL_UNITS:
                raise SqlError(f"unknown interval unit {unit!r}")
            m, d, u = self._INTERVAL_UNITS[key]
            months += int(qty * m)
            days += int(qty * d)
            micros += int(qty * u)
        seen = False
        while True:
            kind, v = self.peek()
            sign = 1
            if kind == "op" and v == "-":
                self.i += 1
                sign, (kind, v) = -1, self.peek()
            if kind == "num":
                self.i += 1
                add(sign * float(v), self.name())
            elif kind == "str":
                self.i += 1
                parts = v.split()
                if len(parts) == 1:
                    add(sign * float(parts[0]), self.name())
                else:
                    for qty, unit in zip(parts[0::2], parts[1::2]):
                        add(sign * float(qty), unit)
            else:
                break
            seen = True
            nxt = self.peek()
            if not (nxt[0] in ("num", "str") or nxt == ("op", "-") and self.peek(1)[0] == "num" and
                    self.peek(2)[0] == "id" and self.peek(2)[1].lower().rstrip("s") in self._INTERVAL_UNITS):
                break
        if not seen:
            raise SqlError("INTERVAL needs a quantity and a unit")
        return Column("interval", (months, days, micros))

    @staticmethod
    def _shift(value: Column, interval: Column, sign: int) -> Column:
        months, days, micros = (sign * x for x in interval._args)
        return Column("call", ("add_interval", (value, months, days, micros)))

    def _postfix(self, col: Column) -> Column:
        """``expr[index]`` / ``expr['key']`` and ``expr.field`` after a call or parenthesis."""
        while True:
            if self.op("["):
                key = self.expr()
                self.expect_op("]")
                col = col.getItem(key._args[0] if key._kind == "lit" else key)
            elif col._kind != "col" and self.peek() == ("op", ".") and self.peek(1)[0] in ("id", "qid"):
                self.i += 1
                col = col.getField(self.name())
            else:
                return col

    def _type_text(self) -> str:
        """The type after ``CAST(x AS``: everything up to the parenthesis that closes the CAST."""
        parts, depth = [], 0
        while True:
            kind, v = self.peek()
            if kind == "eof":
                raise SqlError("unterminated CAST")
            if kind == "op" and v in ("(", "<"):
                depth += 1
            elif kind == "op" and v in (")", ">"):
                if depth == 0 and v == ")":
                    break
                depth -= 1
            parts.append(v)
            self.i += 1
        return "".join(parts)

    def _case(self) -> Column:
        self.kw("case")
        out: Optional[Column] = None
        while self.kw("when"):
            cond = self.expr()
            if not self.kw("then"):
                raise SqlError("CASE WHEN needs THEN")
            value = self.expr()
            out = F.when(cond, value) if out is None else out.when(cond, value)
        if out is None:
            raise SqlError("CASE without WHEN")
        if self.kw("else"):
            out = out.otherwise(self.expr())
        if not self.kw("end"):
            raise SqlError("CASE needs END")
        return out

    def _call(self) -> Column:
        fn_name = self.name()
        low = fn_name.lower()
        self.expect_op("(")
        if low == "count" and self.op("*"):
            self.expect_op(")")
            return F.count("*")
        distinct = self.kw("distinct")
        args: List[Column] = []
        if not self.op(")"):
            args.append(self.expr())
            while self.op(","):
                args.append(self.expr())
            self.expect_op(")")
        if low == "count" and distinct:
            return F.countDistinct(*args)
        alias = {"avg": "avg", "mean": "mean", "power": "pow", "ln": "log", "ceiling": "ceil", "ucase": "upper",
                 "lcase": "lower", "ifnull": "coalesce", "nvl": "coalesce", "std": "stddev"}
        target = alias.get(low, low)
        fn = getattr(F, target, None)
        if fn is None:          # camelCase names such as countDistinct, dayofmonth variants
            for cand in dir(F):
                if cand.lower() == target and callable(getattr(F, cand)):
                    fn = getattr(F, cand)
                    break
        if fn is None or not callable(fn):
            udf = getattr(self.session, "_registered_udfs", {}).get(fn_name) if self.session is not None else None
            if udf is None:
                raise SqlError(f"unknown SQL function {fn_name!r}")
            fn = udf
        out = fn(*args)
        if self.kw("over"):
            out = out.over(self._window_spec())
        return out

    def _window_spec(self):
        from raydp_b200.sql.window import Window, WindowSpec
        if not self.op("("):
            name = self.name()                              # OVER w  (WINDOW w AS (...))
            spec = getattr(self, "named_windows", {}).get(name.lower())
            if spec is None:
                raise SqlError(f"unknown window {name!r}")
            return spec
        spec = WindowSpec()
        if self.kw("partition", "by") or self.kw("distribute", "by"):
            parts = [self.expr()]
            while self.op(","):
                parts.append(self.expr())
            spec = spec.partitionBy(*parts)
        if self.kw("order", "by") or self.kw("sort", "by"):
            spec = spec.orderBy(*self._order_items())
        for kind in ("rows", "range"):
            if self.kw(kind):
                def bound() -> int:
                    if self.kw("unbounded", "preceding"):
                        return Window.unboundedPreceding
                    if self.kw("unbounded", "following"):
                        return Window.unboundedFollowing
                    if self.kw("current", "row"):
                        return 0
                    n = int(self._literal_value())
                    if self.kw("preceding"):
                        return -n
                    if self.kw("following"):
                        return n
                    raise SqlError("frame bound needs PRECEDING or FOLLOWING")
                if self.kw("between"):
                    lo = bound()
                    if not self.kw("and"):
                        raise SqlError("frame BETWEEN needs AND")
                    hi = bound()
                else:
                    lo, hi = bound(), 0
                spec = spec.rowsBetween(lo, hi) if kind == "rows" else spec.rangeBetween(lo, hi)
        self.expect_op(")")
        return spec

    def _order_items(self) -> List[Column]:
        out = []
        while True:
            e = self.expr()
            direction = "ascending"
            if self.kw("desc"):
                direction = "descending"
            else:
                self.kw("asc")
            placement = None
            if self.kw("nulls", "first"):
                placement = "at_start"
            elif self.kw("nulls", "last"):
                placement = "at_end"
            if placement is None:
                out.append(e.desc() if direction == "descending" else e.asc())
            else:
                out.append(Column("sort", (e, direction, placement)))
            if not self.op(","):
                return out


def _is_aggregate(col: Column) -> bool:
    """True if the expression tree contains an aggregate call."""
    if getattr(col, "_kind", None) == "agg":
        return True
    if getattr(col, "_kind", None) == "window":      # sum(x) OVER (...) is not a grouping aggregate
        return False
    for a in getattr(col, "_args", ()) or ():
        if isinstance(a, Column) and _is_aggregate(a):
            return True
        if isinstance(a, (tuple, list)) and any(isinstance(x, Column) and _is_aggregate(x) for x in a):
            return True
    return False


def _signature(node) -> str:
    """Structural fingerprint of an expression (``Column.__eq__`` builds an expression, it does not compare)."""
    if isinstance(node, Column):
        return f"{node._kind}({', '.join(_signature(a) for a in node._args)})"
    if isinstance(node, (tuple, list)):
        return "[" + ", ".join(_signature(a) for a in node) + "]"
    return repr(node)


def _replace_aggregates(expr: Column, registry: Dict[str, str], hidden: List[Column]) -> Column:
    """Aggregates inside HAVING / ORDER BY become references to aggregate output columns; an aggregate the
    SELECT list does not compute is added as a hidden output."""
    if not isinstance(expr, Column):
        return expr
    if expr._kind == "agg":
        sig = _signature(expr)
        if sig not in registry:
            registry[sig] = f"__hidden_agg_{len(hidden)}"
            hidden.append(expr.alias(registry[sig]))
        return F.col(registry[sig])
    if expr._kind in ("col", "lit"):
        return expr

    def walk(node):
        if isinstance(node, Column):
            return _replace_aggregates(node, registry, hidden)
        if isinstance(node, tuple):
            return tuple(walk(x) for x in node)
        if isinstance(node, list):
            return [walk(x) for x in node]
        return node
    return Column(expr._kind, tuple(walk(a) for a in expr._args), expr._name)


def parse_expression(text: str) -> Column:
    """One SQL expression (optionally ``... AS alias``) -> :class:`Column`."""
    p = _Parser(tokenize(text))
    col = p.expr()
    if p.kw("as"):
        col = col.alias(p.name())
    elif not p.done() and p.peek()[0] in ("id", "qid") and p.peek()[1].lower() not in _KEYWORDS:
        col = col.alias(p.name())
    if not p.done():
        raise SqlError(f"unexpected trailing input {p.peek()[1]!r} in expression {text!r}")
    return col


# --------------------------------------------------------------------------------------------------
# catalog: temp views (per session) + parquet-backed tables under spark.sql.warehouse.dir
# --------------------------------------------------------------------------------------------------
class Catalog:
    """``spark.catalog``: temporary views live in the session; tables are parquet directories under the
    warehouse (``<warehouse>/<db>.db/<table>``, database ``default`` -> ``<warehouse>/<table>``), which is
    what ``enable_hive`` sessions persist to in place of a Hive metastore."""

    def __init__(self, session):
        self._session = session
        self._views = {}
        self._current_db = "default"

    # ---- locations
    def warehouse(self) -> str:
        import os
        path = self._session.conf.get("spark.sql.warehouse.dir", None) or os.path.join(os.getcwd(), "spark-warehouse")
        return path[len("file://"):] if path.startswith("file://") else path

    def _table_dir(self, name: str) -> str:
        import os
        db, _, table = name.rpartition(".")
        db = db or self._current_db
        base = self.warehouse() if db == "default" else os.path.join(self.warehouse(), db + ".db")
        return os.path.join(base, table)

    # ---- pyspark surface
    def currentDatabase(self) -> str:
        return self._current_db

    def setCurrentDatabase(self, db: str) -> None:
        self._current_db = db

    def listDatabases(self) -> List[str]:
        import os
        dbs = {"default"}
        if os.path.isdir(self.warehouse()):
            dbs.update(d[:-3] for d in os.listdir(self.warehouse()) if d.endswith(".db"))
        return sorted(dbs)

    def createDatabase(self, db: str) -> None:
        import os
        if db != "default":
            os.makedirs(os.path.join(self.warehouse(), db + ".db"), exist_ok=True)

    def listTables(self, dbName: Optional[str] = None) -> List[Tuple[str, str, bool]]:
        """``[(name, database, isTemporary)]``"""
        import os
        db = dbName or self._current_db
        base = self.warehouse() if db == "default" else os.path.join(self.warehouse(), db + ".db")
        out = []
        if os.path.isdir(base):
            out = [(t, db, False) for t in sorted(os.listdir(base))
                   if os.path.isdir(os.path.join(base, t)) and not t.endswith(".db")]
        return out + [(v, "", True) for v in sorted(self._views)]

    def tableExists(self, name: str) -> bool:
        import os
        return name in self._views or os.path.isdir(self._table_dir(name))

    def dropTempView(self, name: str) -> bool:
        return self._views.pop(name, None) is not None

    def register_view(self, name: str, df) -> None:
        self._views[name] = df

    def lookup(self, name: str):
        import os
        if name in self._views:
            return self._views[name]
        path = self._table_dir(name)
        if os.path.isdir(path):
            return self._session.read.parquet(path)
        raise SqlError(f"Table or view not found: {name}")

    # ---- the rest of ``spark.catalog``
    def databaseExists(self, dbName: str) -> bool:
        return dbName in self.listDatabases()

    def getDatabase(self, dbName: str):
        import collections
        import os
        if not self.databaseExists(dbName):
            raise SqlError(f"Database '{dbName}' not found")
        Database = collections.namedtuple("Database", "name catalog description locationUri")
        path = self.warehouse() if dbName == "default" else os.path.join(self.warehouse(), dbName + ".db")
        return Database(dbName, "spark_catalog", "", "file:" + path)

    def getTable(self, tableName: str):
        import collections
        if not self.tableExists(tableName):
            raise SqlError(f"Table or view not found: {tableName}")
        Table = collections.namedtuple("Table", "name catalog namespace description tableType isTemporary")
        temp = tableName in self._views
        db = None if temp else (tableName.split(".")[0] if "." in tableName else self._current_db)
        return Table(tableName.split(".")[-1], None if temp else "spark_catalog", None if temp else [db], None,
                     "TEMPORARY" if temp else "MANAGED", temp)

    def listColumns(self, tableName: str, dbName: Optional[str] = None):
        import collections
        Col = collections.namedtuple("Column", "name description dataType nullable isPartition isBucket")
        df = self.lookup(f"{dbName}.{tableName}" if dbName else tableName)
        return [Col(f.name, None, f.dataType.simpleString(), f.nullable, False, False) for f in df.schema.fields]

    def listFunctions(self, dbName: Optional[str] = None, pattern: Optional[str] = None):
        import collections
        import fnmatch
        Fn = collections.namedtuple("Function", "name catalog namespace description className isTemporary")
        names = sorted(n for n in dir(F) if not n.startswith("_") and callable(getattr(F, n)))
        names += sorted(set(getattr(self._session, "_registered_udfs", {})))
        if pattern:
            names = [n for n in names if fnmatch.fnmatch(n, pattern)]
        return [Fn(n, None, None, None, "", n in getattr(self._session, "_registered_udfs", {})) for n in names]

    def functionExists(self, functionName: str, dbName: Optional[str] = None) -> bool:
        return any(f.name.lower() == functionName.lower() for f in self.listFunctions())

    def dropGlobalTempView(self, name: str) -> bool:
        gone = self._views.pop(f"global_temp.{name}", None) is not None
        return (self._views.pop(name, None) is not None) or gone

    def cacheTable(self, tableName: str, storageLevel=None) -> None:
        self.lookup(tableName).cache()

    def uncacheTable(self, tableName: str) -> None:
        self.lookup(tableName).unpersist()

    def isCached(self, tableName: str) -> bool:
        return bool(getattr(self.lookup(tableName), "is_cached", False))

    def clearCache(self) -> None:
        for df in self._views.values():
            df.unpersist()

    def refreshTable(self, tableName: str) -> None:
        pass                                            # tables are read from their files on every lookup

    def refreshByPath(self, path: str) -> None:
        pass

    def recoverPartitions(self, tableName: str) -> None:
        pass

    def currentCatalog(self) -> str:
        return "spark_catalog"

    def listCatalogs(self, pattern: Optional[str] = None):
        import collections
        return [collections.namedtuple("CatalogMetadata", "name description")("spark_catalog", None)]

    def createTable(self, tableName: str, path: Optional[str] = None, source: Optional[str] = None, schema=None,
                    description: Optional[str] = None, **options):
        """An external table over files at ``path`` (registered as a view of the reader), or an empty managed
        table with ``schema``."""
        if path is not None:
            reader = self._session.read.format(source or "parquet").options(**options)
            if schema is not None:
                reader = reader.schema(schema)
            df = reader.load(path)
            self._views[tableName] = df
            return df
        if schema is None:
            raise SqlError("createTable needs a path or a schema")
        empty = self._session.createDataFrame([], schema)
        self.save_table(empty, tableName, "errorifexists")
        return self.lookup(tableName)

    createExternalTable = createTable

    def save_table(self, df, name: str, mode: str = "errorifexists") -> None:
        import os
        target = self._table_dir(name)
        if mode == "overwrite" and os.path.isdir(target):
            # the new contents may be computed FROM the table (INSERT OVERWRITE t SELECT ... FROM t, TRUNCATE):
            # evaluate them before the old files go away
            df = df._session._from_tables(df._run("collect") or [df._arrow_schema().empty_table()])
        df.write.mode(mode).parquet(target)

    def drop_table(self, name: str, if_exists: bool = False) -> None:
        import os
        import shutil
        path = self._table_dir(name)
        if os.path.isdir(path):
            shutil.rmtree(path)
        elif not if_exists and name not in self._views:
            raise SqlError(f"Table or view not found: {name}")
        self._views.pop(name, None)


# --------------------------------------------------------------------------------------------------
# statements
# --------------------------------------------------------------------------------------------------
def _qualified_name(p: _Parser) -> str:
    name = p.name()
    while p.op("."):
        name += "." + p.name()
    return name


_JOIN_WORDS = {"join", "inner", "left", "right", "full", "cross", "on", "using", "outer", "semi", "anti", "natural"}
_CLAUSE_WORDS = {"where", "group", "having", "order", "limit", "union", "intersect", "except", "minus", "window",
                 "offset"}


def _parse_select(p: _Parser, session):
    """Kept name of the statement entry point: a full query (WITH / set operations / ORDER BY / LIMIT)."""
    p.session = session
    return _parse_query(p, session)


def _parse_query(p: _Parser, session):
    """``[WITH name AS (query), ...] core {UNION [ALL] | INTERSECT | EXCEPT core} [ORDER BY ...] [LIMIT n]``"""
    p.session = session
    saved_ctes = dict(p.ctes)
    if p.kw("with"):
        while True:
            name = p.name()
            p.kw("as")
            p.expect_op("(")
            p.ctes[name.lower()] = p._subquery()
            if not p.op(","):
                break
    df = _parse_term(p, session)
    combined = False
    while True:
        if p.kw("union"):
            keep_all = p.kw("all")
            p.kw("distinct")
            right = _parse_term(p, session, tail=False)
            df = df.union(right) if keep_all else df.union(right).distinct()
        elif p.kw("intersect"):
            keep_all = p.kw("all")
            right = _parse_term(p, session, tail=False)
            df = df.intersectAll(right) if keep_all else df.intersect(right)
        elif p.kw("except") or p.kw("minus"):
            keep_all = p.kw("all")
            right = _parse_term(p, session, tail=False)
            df = df.exceptAll(right) if keep_all else df.subtract(right)
        else:
            break
        combined = True
    if combined:                                    # the tail belongs to the combined result
        if p.kw("order", "by"):
            df = df.orderBy(*p._order_items())
        if p.kw("limit"):
            df = df.limit(int(p._literal_value()))
    p.ctes = saved_ctes
    return df


def _parse_term(p: _Parser, session, tail: bool = True):
    if p.peek() == ("op", "(") and p.peek(1)[0] == "id" and p.peek(1)[1].lower() in ("select", "with"):
        p.i += 1
        return p._subquery()
    return _parse_core(p, session, tail)


def _table_ref(p: _Parser, session) -> Tuple[Any, Optional[str]]:
    """``name [AS alias]`` or ``( query ) [AS] alias`` -> (DataFrame, alias)."""
    if p.op("("):
        df = p._subquery()
        default_alias = None
    else:
        name = _qualified_name(p)
        df = p.ctes.get(name.lower())
        if df is None:
            df = session.catalog.lookup(name)
        default_alias = name.split(".")[-1]
    alias = default_alias
    if p.kw("as"):
        alias = p.name()
    elif p.peek()[0] in ("id", "qid") and p.peek()[1].lower() not in _KEYWORDS | _JOIN_WORDS | _CLAUSE_WORDS:
        alias = p.name()
    return df, alias


def _side_of(name: str, left_aliases: set, right_alias: Optional[str], left_cols, right_cols) -> Tuple[str, str]:
    """``alias.col`` / ``col`` of an ON equality -> (``"l"`` | ``"r"``, bare column name)."""
    if "." in name:
        qual, bare = name.rsplit(".", 1)
        if qual == right_alias and bare in right_cols:
            return "r", bare
        if qual in left_aliases and (bare in left_cols or f"{qual}.{bare}" in left_cols):
            return "l", bare if bare in left_cols else f"{qual}.{bare}"
        name = bare
    if name in left_cols and name not in right_cols:
        return "l", name
    if name in right_cols and name not in left_cols:
        return "r", name
    return ("l", name) if name in left_cols else ("r", name)


def _parse_core(p: _Parser, session, tail: bool = True):
    if not p.kw("select"):
        raise SqlError("expected SELECT")
    distinct = p.kw("distinct")
    p.kw("all")
    items: List[Any] = []          # "*" or Column
    while True:
        if p.op("*"):
            items.append("*")
        else:
            col = p.expr()
            if p.kw("as"):
                col = col.alias(p.name())
            elif p.peek()[0] in ("id", "qid") and p.peek()[1].lower() not in _KEYWORDS | _CLAUSE_WORDS:
                col = col.alias(p.name())
            elif col._kind == "col" and "." in col._args[0] and not col._args[0].endswith(".*") and col._name is None:
                col = col.alias(col._args[0].rsplit(".", 1)[1])        # a.k shows up as k
            items.append(col)
        if not p.op(","):
            break
    if not p.kw("from"):
        # SELECT <constant expressions>: one row
        one = session.range(1)
        return one.select(*[c for c in items if not isinstance(c, str)])
    df, alias = _table_ref(p, session)
    left_aliases = {alias} if alias else set()
    while True:                                     # [INNER | LEFT | RIGHT | FULL [OUTER] | CROSS | LEFT SEMI/ANTI] JOIN
        how = None
        natural = p.kw("natural")
        if p.kw("join") or p.kw("inner", "join"):
            how = "inner"
        elif p.kw("cross", "join"):
            how = "cross"
        elif p.op(","):
            how = "cross"                           # FROM a, b
        else:
            for side in ("left", "right", "full"):
                if p.at_kw(side):
                    p.i += 1
                    if p.kw("semi"):
                        how = side + "_semi"
                    elif p.kw("anti"):
                        how = side + "_anti"
                    else:
                        p.kw("outer")
                        how = side
                    if not p.kw("join"):
                        raise SqlError(f"expected JOIN after {side.upper()}")
                    break
        if how is None:
            break
        other, other_alias = _table_ref(p, session)
        # columns both sides have (other than the join keys, known below) stay reachable as <alias>.<column>
        pairs: List[Tuple[str, str]] = []
        using: Optional[List[str]] = None
        if natural:
            using = [c for c in df.columns if c in other.columns]
        elif how == "cross":
            using = None
        elif p.kw("using"):
            p.expect_op("(")
            using = [p.name()]
            while p.op(","):
                using.append(p.name())
            p.expect_op(")")
        elif p.kw("on"):
            from raydp_b200.sql.dataframe import _equi_join_pairs
            for a, b in _equi_join_pairs(p.expr()):
                side_a, name_a = _side_of(a, left_aliases, other_alias, df.columns, other.columns)
                side_b, name_b = _side_of(b, left_aliases, other_alias, df.columns, other.columns)
                if side_a == side_b:
                    raise SqlError(f"join condition {a} = {b} does not relate the two sides")
                pairs.append((name_a, name_b) if side_a == "l" else (name_b, name_a))
        else:
            raise SqlError("JOIN needs ON <equalities> or USING (<columns>)")
        key_right = set(using or []) | {r for _, r in pairs}
        clashes = [c for c in other.columns if c in df.columns and c not in key_right]
        if clashes and how not in ("left_semi", "left_anti"):
            tag = other_alias or "right"
            for c in clashes:
                other = other.withColumnRenamed(c, f"{tag}.{c}")
        if how == "cross":
            df = df.crossJoin(other)
        elif using is not None:
            df = df.join(other, using, how)
        else:
            left_keys, right_keys = [l for l, _ in pairs], [r for _, r in pairs]
            renamed = other
            for lk, rk in pairs:
                if lk != rk:
                    if lk in renamed.columns:       # the left key's name is taken on the right by another column
                        renamed = renamed.withColumnRenamed(lk, f"{other_alias or 'right'}.{lk}")
                    renamed = renamed.withColumnRenamed(rk, lk)
            df = df.join(renamed, left_keys, how)
        if other_alias:
            left_aliases.add(other_alias)
    if p.kw("where"):
        df = df.filter(p.expr())
    key_exprs: List[Column] = []
    if p.kw("group", "by"):
        key_exprs.append(p.expr())
        while p.op(","):
            key_exprs.append(p.expr())
    keys: List[str] = [e._output_name() for e in key_exprs]
    having = p.expr() if p.kw("having") else None
    order: List[Column] = []
    limit = None
    if tail:
        if p.kw("order", "by") or p.kw("sort", "by"):
            order = p._order_items()
        if p.kw("limit"):
            limit = int(p._literal_value())

    def order_parts(item: Column) -> Tuple[Column, str, Optional[str]]:
        return item._args[0], item._args[1], (item._args[2] if len(item._args) > 2 else None)

    def rebuild(expr: Column, direction: str, placement: Optional[str]) -> Column:
        return Column("sort", (expr, direction) + ((placement,) if placement else ()))

    cols = [c for c in items if not isinstance(c, str)]
    # ORDER BY 1, 2: ordinals of the select list
    for i, item in enumerate(order):
        e, direction, placement = order_parts(item)
        if e._kind == "lit" and isinstance(e._args[0], int) and 1 <= e._args[0] <= len(cols):
            order[i] = rebuild(F.col(cols[e._args[0] - 1]._output_name()), direction, placement)
    grouped = bool(keys) or any(_is_aggregate(c) for c in cols) or (having is not None and _is_aggregate(having))
    if grouped:
        from raydp_b200.sql.dataframe import _agg_default_name
        select_sigs = {_signature(Column(c._kind, c._args)): c for c in cols if not _is_aggregate(c)}
        # GROUP BY 1, 2 (ordinals), GROUP BY <select alias>, GROUP BY <expression that a select item repeats>
        for i, e in enumerate(key_exprs):
            if e._kind == "lit" and isinstance(e._args[0], int) and 1 <= e._args[0] <= len(cols):
                key_exprs[i] = cols[e._args[0] - 1]
            elif e._kind == "col" and e._args[0] not in df.columns:
                match = [c for c in cols if c._name == e._args[0]]
                if match:
                    key_exprs[i] = match[0]
            elif e._name is None and _signature(e) in select_sigs:
                key_exprs[i] = select_sigs[_signature(e)]
        keys = [e._output_name() for e in key_exprs]
        keys = [k.split(".", 1)[1] if k not in df.columns and "." in k and k.split(".", 1)[1] in df.columns else k
                for k in keys]
        key_sigs = {_signature(Column(e._kind, e._args)): e._output_name() for e in key_exprs}
        agg_items: List[Column] = []
        registry: Dict[str, str] = {}
        post: List[Column] = []
        for c in cols:
            bare_sig = _signature(Column(c._kind, c._args))
            if _is_aggregate(c):
                name = c._name or _agg_default_name(c)
                agg_items.append(c.alias(name))
                if c._kind == "agg":
                    registry[_signature(Column(c._kind, c._args))] = name
                post.append(F.col(name))
            elif c._output_name() in keys:
                post.append(F.col(c._output_name()))
            elif bare_sig in key_sigs:
                post.append(F.col(key_sigs[bare_sig]).alias(c._output_name()))
            else:
                post.append(c)                      # an expression over the grouping keys: applied afterwards
        if any(isinstance(c, str) for c in items):
            raise SqlError("SELECT * cannot be combined with GROUP BY")
        hidden: List[Column] = []
        if having is not None:
            having = _replace_aggregates(having, registry, hidden)
        order = [rebuild(_replace_aggregates(order_parts(o)[0], registry, hidden), *order_parts(o)[1:]) for o in order]
        if agg_items or hidden:
            out = df.groupBy(*key_exprs).agg(*(agg_items + hidden))
        else:
            out = df.groupBy(*key_exprs).agg(F.count("*").alias("__n")).drop("__n")
        if having is not None:
            out = out.filter(having)
        out_names = [c._output_name() for c in post]
        if order and not all(order_parts(o)[0]._kind == "col" and order_parts(o)[0]._output_name() in out_names
                             for o in order):
            resolved = []
            for o in order:
                e, direction, placement = order_parts(o)
                alias_hit = [c for c in post if c._name == e._output_name()] if e._kind == "col" and \
                    e._args[0] not in out.columns else []
                resolved.append(rebuild(alias_hit[0] if alias_hit else e, direction, placement))
            out = out.orderBy(*resolved)
            order = []
        out = out.select(*post)
    else:
        out = df
        proj: List[Any] = []
        for c in items:
            proj.extend(df.columns if isinstance(c, str) else [c])
        out_names = [c if isinstance(c, str) else c._output_name() for c in proj]
        # ORDER BY on output names (aliases) sorts after the projection; anything else is evaluated on
        # the input first (the global sort's order survives the narrow projection that follows)
        if order and not all(order_parts(o)[0]._kind == "col" and order_parts(o)[0]._output_name() in out_names
                             for o in order):
            out = out.orderBy(*order)
            order = []
        star_only = len(items) == 1 and isinstance(items[0], str)
        if not star_only or out.columns != df.columns:
            out = out.select(*proj)
    if distinct:
        out = out.distinct()
    if order:
        out = out.orderBy(*order)
    if limit is not None:
        out = out.limit(limit)
    return out


def run_sql(session, query: str):
    """Execute one statement; always returns a DataFrame (empty for DDL)."""
    p = _Parser(tokenize(query), session)
    cat = session.catalog
    if p.at_kw("select") or p.at_kw("with") or p.peek() == ("op", "("):
        out = _parse_select(p, session)
    elif p.kw("show", "tables"):
        db = None
        if p.kw("in") or p.kw("from"):
            db = p.name()
        rows = [(db_ or "", name, temp) for name, db_, temp in cat.listTables(db)]
        return session.createDataFrame(rows, ["namespace", "tableName", "isTemporary"]) if rows else \
            session.createDataFrame([("", "", False)], ["namespace", "tableName", "isTemporary"]).limit(0)
    elif p.kw("show", "databases"):
        return session.createDataFrame([(d,) for d in cat.listDatabases()], ["namespace"])
    elif p.kw("create", "database") or p.kw("create", "schema"):
        p.kw("if", "not", "exists")
        cat.createDatabase(p.name())
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    elif p.kw("use"):
        cat.setCurrentDatabase(p.name())
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    elif p.kw("drop", "table") or p.kw("drop", "view"):
        if_exists = p.kw("if", "exists")
        cat.drop_table(_qualified_name(p), if_exists)
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    elif p.kw("create", "table") or p.kw("create", "or", "replace", "table") or p.kw("create", "external", "table"):
        if_not_exists = p.kw("if", "not", "exists")
        name = _qualified_name(p)
        schema = None
        if p.peek() == ("op", "("):                          # CREATE TABLE t (a int, b string)
            p.i += 1
            depth, parts = 0, []
            while not (p.peek() == ("op", ")") and depth == 0):
                kind, v = p.peek()
                if kind == "eof":
                    raise SqlError("unterminated column list")
                depth += (v in ("(", "<")) - (v in (")", ">")) if kind == "op" else 0
                parts.append("," if (kind, v) == ("op", ",") else v)
                p.i += 1
            p.i += 1
            from raydp_b200.sql.types import parse_ddl
            schema = parse_ddl(" ".join(parts).replace(" ,", ",").replace("< ", "<").replace(" >", ">")
                               .replace(" <", "<").replace("( ", "(").replace(" )", ")").replace(" (", "("))
        if p.kw("using"):
            p.name()
        while p.kw("partitioned", "by") or p.kw("location") or p.kw("comment") or p.kw("tblproperties"):
            if p.peek()[0] == "str":
                p.i += 1
            elif p.op("("):
                while not p.op(")"):
                    p.i += 1
        if p.kw("as"):
            cat.save_table(_parse_select(p, session), name, "overwrite")
        elif schema is not None:
            if not (if_not_exists and cat.tableExists(name)):
                cat.save_table(session.createDataFrame([], schema), name, "errorifexists")
        else:
            raise SqlError("CREATE TABLE needs a column list or AS SELECT")
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    elif p.kw("create", "view") or p.kw("create", "or", "replace", "view") or p.kw("create", "temp", "view") or \
            p.kw("create", "temporary", "view") or p.kw("create", "or", "replace", "temp", "view") or \
            p.kw("create", "or", "replace", "temporary", "view") or p.kw("create", "global", "temp", "view") or \
            p.kw("create", "or", "replace", "global", "temp", "view"):
        p.kw("if", "not", "exists")
        name = _qualified_name(p)
        if not p.kw("as"):
            raise SqlError("CREATE VIEW needs AS SELECT")
        cat.register_view(name, _parse_select(p, session))
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    elif p.kw("insert", "into") or p.kw("insert", "overwrite", "table") or p.kw("insert", "overwrite"):
        overwrite = p.t[1][1].lower() == "overwrite"
        p.kw("table")
        name = _qualified_name(p)
        if p.kw("values"):
            target = cat.lookup(name)
            rows = []
            while True:
                p.expect_op("(")
                row = [p.expr()]
                while p.op(","):
                    row.append(p.expr())
                p.expect_op(")")
                rows.append(tuple(session.range(1).select(*[c.alias(f"c{i}") for i, c in enumerate(row)]).collect()[0]))
                if not p.op(","):
                    break
            data = session.createDataFrame(rows, target.schema)
        else:
            data = _parse_select(p, session)
            if cat.tableExists(name):                    # columns are matched by position, typed by the table
                target = cat.lookup(name)
                if len(target.columns) != len(data.columns):
                    raise SqlError(f"INSERT INTO {name}: the table has {len(target.columns)} columns, the query "
                                   f"returns {len(data.columns)}")
                data = data.toDF(*target.columns)
                data = data.select(*[F.col(f.name).cast(f.dataType) for f in target.schema.fields])
        cat.save_table(data, name, "overwrite" if overwrite else "append")
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    elif p.kw("truncate", "table"):
        name = _qualified_name(p)
        cat.save_table(cat.lookup(name).limit(0), name, "overwrite")
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    elif p.kw("describe") or p.kw("desc"):
        p.kw("table")
        p.kw("extended")
        df = cat.lookup(_qualified_name(p))
        return session.createDataFrame([(f.name, f.dataType.simpleString(), None) for f in df.schema.fields],
                                       "col_name string, data_type string, comment string")
    elif p.kw("show", "columns"):
        if not (p.kw("in") or p.kw("from")):
            raise SqlError("SHOW COLUMNS needs IN <table>")
        df = cat.lookup(_qualified_name(p))
        return session.createDataFrame([(c,) for c in df.columns], ["col_name"])
    elif p.kw("show", "functions"):
        return session.createDataFrame([(f.name,) for f in cat.listFunctions()], ["function"])
    elif p.kw("show", "views"):
        return session.createDataFrame([("", v, True) for v in sorted(cat._views)] or [("", "", True)],
                                       ["namespace", "viewName", "isTemporary"]).limit(len(cat._views))
    elif p.kw("cache", "table") or p.kw("cache", "lazy", "table"):
        cat.cacheTable(_qualified_name(p))
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    elif p.kw("uncache", "table"):
        p.kw("if", "exists")
        cat.uncacheTable(_qualified_name(p))
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    elif p.kw("clear", "cache"):
        cat.clearCache()
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    elif p.kw("refresh"):
        p.kw("table")
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    elif p.kw("set"):
        text = query.strip().rstrip(";")[3:].strip()
        if not text:
            return session.createDataFrame(session._conf.getAll() or [("", "")], ["key", "value"])
        key, sep, value = text.partition("=")
        if sep:
            session.conf.set(key.strip(), value.strip())
            return session.createDataFrame([(key.strip(), value.strip())], ["key", "value"])
        return session.createDataFrame([(key.strip(), session.conf.get(key.strip(), "<undefined>"))], ["key", "value"])
    elif p.kw("explain"):
        p.kw("extended")
        _parse_select(p, session).explain()
        return session.createDataFrame([("see stdout",)], ["plan"])
    elif p.kw("values"):
        rows = []
        while True:
            p.expect_op("(")
            row = [p.expr()]
            while p.op(","):
                row.append(p.expr())
            p.expect_op(")")
            rows.append(tuple(session.range(1).select(*[c.alias(f"col{i + 1}") for i, c in enumerate(row)]).collect()[0]))
            if not p.op(","):
                break
        return session.createDataFrame(rows, [f"col{i + 1}" for i in range(len(rows[0]))])
    else:
        raise SqlError(f"unsupported SQL statement: {query.strip()[:60]!r} (see raydp_b200/sql/sqlparser.py)")
    if not p.done():
        raise SqlError(f"unexpected trailing input {p.peek()[1]!r}")
    return out
