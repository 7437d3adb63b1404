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
    | (?P<op><=|>=|<>|!=|==|[-+*/%(),.<>=])
    )""", re.X)

_AGGREGATES = {"count", "sum", "avg", "mean", "min", "max", "first", "stddev", "variance", "countdistinct"}
_KEYWORDS = {"select", "from", "where", "group", "by", "having", "order", "limit", "as", "and", "or", "not",
             "is", "null", "in", "between", "case", "when", "then", "else", "end", "cast", "distinct", "asc",
             "desc", "true", "false", "like", "join", "inner", "left", "right", "full", "cross", "on", "using"}


class SqlError(ValueError):
    pass


def tokenize(text: str) -> List[Tuple[str, str]]:
    out, pos = [], 0
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
    def __init__(self, tokens: List[Tuple[str, str]]):
        self.t = tokens
        self.i = 0

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
            elif self.at_kw("not") and self.peek(1)[1].lower() in ("in", "between"):
                self.i += 1
                left = ~self._in_or_between(left)
            elif self.at_kw("in") or self.at_kw("between"):
                left = self._in_or_between(left)
            else:
                kind, v = self.peek()
                if kind == "op" and v in ("=", "==", "<>", "!=", "<", "<=", ">", ">="):
                    self.i += 1
                    right = self.additive()
                    left = {"=": left == right, "==": left == right, "<>": left != right, "!=": left != right,
                            "<": left < right, "<=": left <= right, ">": left > right, ">=": left >= right}[v]
                else:
                    return left

    def _in_or_between(self, left: Column) -> Column:
        if self.kw("in"):
            self.expect_op("(")
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
                left = left + self.multiplicative()
            elif self.op("-"):
                left = left - self.multiplicative()
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
            inner = self.expr()
            self.expect_op(")")
            return inner
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
        if low == "cast" and self.peek(1) == ("op", "("):
            self.i += 2
            inner = self.expr()
            if not self.kw("as"):
                raise SqlError("CAST needs AS <type>")
            type_name = self.name()
            self.expect_op(")")
            return inner.cast(type_name.lower())
        if self.peek(1) == ("op", "("):
            return self._call()
        self.i += 1
        # qualified name  view.column -> column
        while self.peek() == ("op", ".") and self.peek(1)[0] in ("id", "qid"):
            self.i += 1
            v = self.name()
        return F.col(v)

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
            raise SqlError(f"unknown SQL function {fn_name!r}")
        return fn(*args)


def _is_aggregate(col: Column) -> bool:
    """True if the expression tree contains an aggregate call."""
    if getattr(col, "_kind", None) == "agg":
        return True
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

    def save_table(self, df, name: str, mode: str = "errorifexists") -> None:
        df.write.mode(mode).parquet(self._table_dir(name))

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


def _parse_select(p: _Parser, session):
    if not p.kw("select"):
        raise SqlError("expected SELECT")
    distinct = p.kw("distinct")
    items: List[Any] = []          # "*" or Column
    while True:
        if p.op("*"):
            items.append("*")
        else:
            col = p.expr()
            if p.kw("as"):
                col = col.alias(p.name())
            elif p.peek()[0] in ("id", "qid") and p.peek()[1].lower() not in _KEYWORDS:
                col = col.alias(p.name())
            items.append(col)
        if not p.op(","):
            break
    if not p.kw("from"):
        # SELECT <constant expressions>: one row
        one = session.range(1)
        return one.select(*[c for c in items if not isinstance(c, str)])
    df = session.catalog.lookup(_qualified_name(p))
    join_words = {"join", "inner", "left", "right", "full", "cross", "on", "using", "outer", "semi", "anti"}

    def skip_alias():
        if p.peek()[0] in ("id", "qid") and p.peek()[1].lower() not in _KEYWORDS | join_words:
            p.name()                                # table alias: columns are resolved by bare name
    skip_alias()
    while True:                                     # [INNER | LEFT | RIGHT | FULL [OUTER] | CROSS | LEFT SEMI/ANTI] JOIN
        how = None
        if p.kw("join") or p.kw("inner", "join"):
            how = "inner"
        elif p.kw("cross", "join"):
            how = "cross"
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
        other = session.catalog.lookup(_qualified_name(p))
        skip_alias()
        if how == "cross":
            df = df.crossJoin(other)
        elif p.kw("using"):
            p.expect_op("(")
            names = [p.name()]
            while p.op(","):
                names.append(p.name())
            p.expect_op(")")
            df = df.join(other, names, how)
        elif p.kw("on"):
            from raydp_b200.sql.dataframe import _equi_join_pairs
            pairs = [(a, b) if a in df.columns else (b, a) for a, b in _equi_join_pairs(p.expr())]
            left_keys, right_keys = [a for a, _ in pairs], [b for _, b in pairs]
            if left_keys == right_keys:
                df = df.join(other, left_keys, how)
            else:
                renamed = other
                for lk, rk in pairs:
                    if lk != rk:
                        renamed = renamed.withColumnRenamed(rk, lk)
                df = df.join(renamed, left_keys, how)
        else:
            raise SqlError("JOIN needs ON <equalities> or USING (<columns>)")
    if p.kw("where"):
        df = df.filter(p.expr())
    key_exprs: List[Column] = []
    if p.kw("group", "by"):
        key_exprs.append(p.expr())
        while p.op(","):
            key_exprs.append(p.expr())
    keys: List[str] = [e._output_name() for e in key_exprs]
    having = p.expr() if p.kw("having") else None
    order: List[Tuple[Column, bool]] = []
    if p.kw("order", "by"):
        while True:
            e = p.expr()
            asc = True
            if p.kw("desc"):
                asc = False
            else:
                p.kw("asc")
            order.append((e, asc))
            if not p.op(","):
                break
    limit = None
    if p.kw("limit"):
        limit = int(p._literal_value())

    cols = [c for c in items if not isinstance(c, str)]
    grouped = bool(keys) or any(_is_aggregate(c) for c in cols)
    if grouped:
        from raydp_b200.sql.dataframe import _agg_default_name
        # GROUP BY 1, 2 (ordinals) and GROUP BY <select alias>
        for i, e in enumerate(key_exprs):
            if e._kind == "lit" and isinstance(e._args[0], int) and 1 <= e._args[0] <= len(cols):
                key_exprs[i] = cols[e._args[0] - 1]
            elif e._kind == "col" and e._args[0] not in df.columns:
                match = [c for c in cols if c._name == e._args[0]]
                if match:
                    key_exprs[i] = match[0]
        keys = [e._output_name() for e in key_exprs]
        agg_items: List[Column] = []
        registry: Dict[str, str] = {}
        post: List[Column] = []
        for c in cols:
            if _is_aggregate(c):
                name = c._name or _agg_default_name(c)
                agg_items.append(c.alias(name))
                if c._kind == "agg":
                    registry[_signature(c)] = name
                post.append(F.col(name))
            elif c._output_name() in keys:
                post.append(F.col(c._output_name()))
            else:
                post.append(c)                      # an expression over the grouping keys: applied afterwards
        hidden: List[Column] = []
        if having is not None:
            having = _replace_aggregates(having, registry, hidden)
        order = [(_replace_aggregates(e, registry, hidden), a) for e, a in order]
        if agg_items or hidden:
            out = df.groupBy(*key_exprs).agg(*(agg_items + hidden))
        else:
            out = df.groupBy(*key_exprs).agg(F.count("*").alias("__n")).drop("__n")
        if having is not None:
            out = out.filter(having)
        out_names = [c._output_name() for c in post]
        if order and not all(e._kind == "col" and e._output_name() in out_names for e, _ in order):
            key_names = [f"__ok{i}" for i in range(len(order))]
            for n, (e, _) in zip(key_names, order):
                alias_hit = [c for c in post if c._name == e._output_name()] if e._kind == "col" and \
                    e._args[0] not in out.columns else []
                out = out.withColumn(n, alias_hit[0] if alias_hit else e)
            out = out.orderBy(*key_names, ascending=[a for _, a in order])
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
        if order and not all(e._kind == "col" and e._output_name() in out_names for e, _ in order):
            key_names = [f"__ok{i}" for i in range(len(order))]
            for n, (e, _) in zip(key_names, order):
                out = out.withColumn(n, e)
            out = out.orderBy(*key_names, ascending=[a for _, a in order])
            order = []
        star_only = len(items) == 1 and isinstance(items[0], str)
        if not star_only or out.columns != df.columns:
            out = out.select(*proj)
    if distinct:
        out = out.distinct()
    if order:
        out = out.orderBy(*[e._output_name() for e, _ in order], ascending=[a for _, a in order])
    if limit is not None:
        out = out.limit(limit)
    return out


def run_sql(session, query: str):
    """Execute one statement; always returns a DataFrame (empty for DDL)."""
    p = _Parser(tokenize(query))
    cat = session.catalog
    if p.at_kw("select"):
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
    elif p.kw("create", "table") or p.kw("create", "or", "replace", "table"):
        p.kw("if", "not", "exists")
        name = _qualified_name(p)
        if p.kw("using"):
            p.name()
        if not p.kw("as"):
            raise SqlError("only CREATE TABLE ... AS SELECT is supported")
        cat.save_table(_parse_select(p, session), name, "overwrite")
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    elif p.kw("insert", "into") or p.kw("insert", "overwrite", "table"):
        overwrite = p.t[1][1].lower() == "overwrite"
        p.kw("table")
        name = _qualified_name(p)
        cat.save_table(_parse_select(p, session), name, "overwrite" if overwrite else "append")
        return session.createDataFrame([(0,)], ["ok"]).limit(0)
    else:
        raise SqlError(f"unsupported SQL statement: {query.strip()[:60]!r} (see raydp_b200/sql/sqlparser.py)")
    if not p.done():
        raise SqlError(f"unexpected trailing input {p.peek()[1]!r}")
    return out
