"""Spark-SQL-shaped data types over Arrow.

Parity: the pyspark type surface the reference touches -- ``StructType``/``StructField``
schemas iterated for ``field.name`` (reference ``examples/pytorch_nyctaxi.py:66``), explicit
schemas for ``spark.read.schema(...)`` (DLRM notebook cell 13) and Arrow<->Spark schema
conversion in ``ray_dataset_to_spark_dataframe`` (reference ``spark/dataset.py:298-305``).
"""
from __future__ import annotations

from typing import Iterator, List, Optional, Sequence, Union

import pyarrow as pa


class DataType:
    arrow: pa.DataType = pa.null()
    name = "null"

    def simpleString(self) -> str:
        return self.name

    def typeName(self) -> str:
        return self.name

    def __eq__(self, other):
        return type(self) is type(other) and self.__dict__ == other.__dict__

    def __hash__(self):
        return hash(type(self).__name__)

    def __repr__(self):
        return f"{type(self).__name__}()"

    def jsonValue(self):
        return self.simpleString() if type(self).__name__ not in _JSON_NAMES else _JSON_NAMES[type(self).__name__]

    def json(self) -> str:
        import json as _json
        return _json.dumps(self.jsonValue(), separators=(",", ":"), sort_keys=True)


_JSON_NAMES = {"LongType": "long", "IntegerType": "integer", "ShortType": "short", "ByteType": "byte",
               "NullType": "void"}


def _simple(cls_name: str, arrow_type: pa.DataType, simple: str):
    return type(cls_name, (DataType,), {"arrow": arrow_type, "name": simple})


NullType = _simple("NullType", pa.null(), "void")
BooleanType = _simple("BooleanType", pa.bool_(), "boolean")
ByteType = _simple("ByteType", pa.int8(), "tinyint")
ShortType = _simple("ShortType", pa.int16(), "smallint")
IntegerType = _simple("IntegerType", pa.int32(), "int")
LongType = _simple("LongType", pa.int64(), "bigint")
FloatType = _simple("FloatType", pa.float32(), "float")
DoubleType = _simple("DoubleType", pa.float64(), "double")
StringType = _simple("StringType", pa.string(), "string")
BinaryType = _simple("BinaryType", pa.binary(), "binary")
DateType = _simple("DateType", pa.date32(), "date")
TimestampType = _simple("TimestampType", pa.timestamp("us"), "timestamp")


class ArrayType(DataType):
    name = "array"

    def __init__(self, elementType: DataType, containsNull: bool = True):
        self.elementType = elementType
        self.containsNull = containsNull

    @property
    def arrow(self):  # type: ignore[override]
        return pa.list_(self.elementType.arrow)

    def simpleString(self):
        return f"array<{self.elementType.simpleString()}>"

    def jsonValue(self):
        return {"type": "array", "elementType": self.elementType.jsonValue(), "containsNull": self.containsNull}


class MapType(DataType):
    name = "map"

    def __init__(self, keyType: DataType, valueType: DataType, valueContainsNull: bool = True):
        self.keyType = keyType
        self.valueType = valueType
        self.valueContainsNull = valueContainsNull

    @property
    def arrow(self):  # type: ignore[override]
        return pa.map_(self.keyType.arrow, self.valueType.arrow)

    def simpleString(self):
        return f"map<{self.keyType.simpleString()},{self.valueType.simpleString()}>"

    def jsonValue(self):
        return {"type": "map", "keyType": self.keyType.jsonValue(), "valueType": self.valueType.jsonValue(),
                "valueContainsNull": self.valueContainsNull}

    def __repr__(self):
        return f"MapType({self.keyType!r}, {self.valueType!r}, {self.valueContainsNull})"


class DecimalType(DataType):
    name = "decimal"

    def __init__(self, precision: int = 10, scale: int = 0):
        self.precision = precision
        self.scale = scale

    @property
    def arrow(self):  # type: ignore[override]
        return pa.decimal128(self.precision, self.scale)

    def simpleString(self):
        return f"decimal({self.precision},{self.scale})"

    def __repr__(self):
        return f"DecimalType({self.precision},{self.scale})"


class StructField:
    def __init__(self, name: str, dataType: DataType, nullable: bool = True, metadata=None):
        self.name = name
        self.dataType = dataType
        self.nullable = nullable
        self.metadata = metadata or {}

    def __repr__(self):
        return f"StructField('{self.name}', {self.dataType!r}, {self.nullable})"

    def __eq__(self, other):
        return (isinstance(other, StructField) and other.name == self.name
                and other.dataType == self.dataType)


class StructType(DataType):
    name = "struct"

    def __init__(self, fields: Optional[Sequence[StructField]] = None):
        self.fields: List[StructField] = list(fields or [])

    def add(self, field: Union[str, StructField], data_type: Optional[DataType] = None,
            nullable: bool = True) -> "StructType":
        if isinstance(field, StructField):
            self.fields.append(field)
        else:
            self.fields.append(StructField(field, parse_type(data_type) if data_type is not None else StringType(),
                                           nullable))
        return self

    def jsonValue(self):
        return {"type": "struct", "fields": [
            {"name": f.name, "type": f.dataType.jsonValue(), "nullable": f.nullable, "metadata": f.metadata}
            for f in self.fields]}

    @classmethod
    def fromJson(cls, json: dict) -> "StructType":  # noqa: A002
        return cls([StructField(f["name"], _type_from_json(f["type"]), f.get("nullable", True), f.get("metadata"))
                    for f in json["fields"]])

    @property
    def names(self) -> List[str]:
        return [f.name for f in self.fields]

    def fieldNames(self) -> List[str]:
        return self.names

    def __iter__(self) -> Iterator[StructField]:
        return iter(self.fields)

    def __len__(self):
        return len(self.fields)

    def __getitem__(self, key):
        if isinstance(key, str):
            for f in self.fields:
                if f.name == key:
                    return f
            raise KeyError(key)
        return self.fields[key]

    @property
    def arrow(self):  # type: ignore[override]
        return pa.struct([pa.field(f.name, f.dataType.arrow) for f in self.fields])

    def to_arrow_schema(self) -> pa.Schema:
        return pa.schema([pa.field(f.name, f.dataType.arrow, f.nullable) for f in self.fields])

    def simpleString(self):
        inner = ",".join(f"{f.name}:{f.dataType.simpleString()}" for f in self.fields)
        return f"struct<{inner}>"

    def __repr__(self):
        return f"StructType([{', '.join(repr(f) for f in self.fields)}])"


_BY_NAME = {
    "boolean": BooleanType, "bool": BooleanType, "tinyint": ByteType, "byte": ByteType,
    "smallint": ShortType, "short": ShortType, "int": IntegerType, "integer": IntegerType,
    "bigint": LongType, "long": LongType, "float": FloatType, "double": DoubleType,
    "string": StringType, "str": StringType, "binary": BinaryType, "date": DateType,
    "timestamp": TimestampType, "void": NullType, "null": NullType, "timestamp_ntz": TimestampType,
}


def _type_from_json(value) -> DataType:
    if isinstance(value, dict):
        kind = value.get("type")
        if kind == "struct":
            return StructType.fromJson(value)
        if kind == "array":
            return ArrayType(_type_from_json(value["elementType"]), value.get("containsNull", True))
        if kind == "map":
            return MapType(_type_from_json(value["keyType"]), _type_from_json(value["valueType"]),
                           value.get("valueContainsNull", True))
        raise ValueError(f"unsupported type json: {value!r}")
    return parse_type(value)


def parse_type(spec: Union[str, DataType, pa.DataType]) -> DataType:
    """``"int"``, ``IntegerType()`` or an Arrow type -> DataType."""
    if isinstance(spec, DataType):
        return spec
    if isinstance(spec, pa.DataType):
        return from_arrow_type(spec)
    text = str(spec).strip()
    key = text.lower()
    if key in _BY_NAME:
        return _BY_NAME[key]()
    if key.startswith("array<") and key.endswith(">"):
        return ArrayType(parse_type(text[6:-1]))
    if key.startswith("map<") and key.endswith(">"):
        parts = _split_top_level(text[4:-1])
        if len(parts) != 2:
            raise ValueError(f"unsupported data type string: {spec!r}")
        return MapType(parse_type(parts[0]), parse_type(parts[1]))
    if key.startswith("struct<") and key.endswith(">"):
        return parse_ddl(text[7:-1])
    if key.startswith(("decimal", "numeric", "dec")):
        inside = key[key.index("(") + 1:key.rindex(")")] if "(" in key else "10,0"
        nums = [int(x) for x in inside.split(",")]
        return DecimalType(nums[0], nums[1] if len(nums) > 1 else 0)
    if key.startswith(("varchar", "char")):
        return StringType()
    raise ValueError(f"unsupported data type string: {spec!r}")


def _split_top_level(text: str, sep: str = ",") -> List[str]:
    """Split on ``sep`` outside ``<...>`` / ``(...)`` nesting."""
    out, depth, cur = [], 0, []
    for ch in text:
        if ch in "<(":
            depth += 1
        elif ch in ">)":
            depth -= 1
        if ch == sep and depth == 0:
            out.append("".join(cur))
            cur = []
        else:
            cur.append(ch)
    if "".join(cur).strip():
        out.append("".join(cur))
    return out


def parse_ddl(ddl: str) -> "StructType":
    """``"a int, b double"`` / ``"a: int, b: string"`` -> StructType (the DDL strings pyspark accepts for
    ``mapInPandas``, ``createDataFrame(schema=...)`` and ``read.schema(...)``)."""
    fields = []
    for part in _split_top_level(ddl):
        part = part.strip()
        if not part:
            continue
        head = part.split("<", 1)[0]
        if ":" in head:                             # "name: type" (struct<...> notation)
            name, _, tname = part.partition(":")
        else:
            name, _, tname = part.partition(" ")
        tname = tname.strip()
        nullable = True
        if tname.lower().endswith(" not null"):
            tname, nullable = tname[:-9].strip(), False
        if not tname:
            raise ValueError(f"cannot parse schema field {part!r}; expected '<name> <type>'")
        fields.append(StructField(name.strip().strip("`"), parse_type(tname), nullable))
    return StructType(fields)


def from_arrow_type(t: pa.DataType) -> DataType:
    if pa.types.is_boolean(t):
        return BooleanType()
    if pa.types.is_int8(t) or pa.types.is_uint8(t):
        return ByteType()
    if pa.types.is_int16(t) or pa.types.is_uint16(t):
        return ShortType()
    if pa.types.is_int32(t) or pa.types.is_uint32(t):
        return IntegerType()
    if pa.types.is_integer(t):
        return LongType()
    if pa.types.is_float32(t) or pa.types.is_float16(t):
        return FloatType()
    if pa.types.is_floating(t):
        return DoubleType()
    if pa.types.is_string(t) or pa.types.is_large_string(t):
        return StringType()
    if pa.types.is_binary(t) or pa.types.is_large_binary(t):
        return BinaryType()
    if pa.types.is_date(t):
        return DateType()
    if pa.types.is_timestamp(t):
        return TimestampType()
    if pa.types.is_list(t) or pa.types.is_large_list(t):
        return ArrayType(from_arrow_type(t.value_type))
    if pa.types.is_struct(t):
        return StructType([StructField(f.name, from_arrow_type(f.type)) for f in t])
    if pa.types.is_map(t):
        return MapType(from_arrow_type(t.key_type), from_arrow_type(t.item_type))
    if pa.types.is_decimal(t):
        return DecimalType(t.precision, t.scale)
    if pa.types.is_fixed_size_list(t):
        return ArrayType(from_arrow_type(t.value_type))
    if pa.types.is_null(t):
        return NullType()
    if pa.types.is_dictionary(t):
        return from_arrow_type(t.value_type)
    raise TypeError(f"Arrow type {t} has no Spark SQL equivalent")


def from_arrow_schema(schema: pa.Schema) -> StructType:
    return StructType([StructField(f.name, from_arrow_type(f.type), f.nullable) for f in schema])


class Row(tuple):
    """Immutable row with attribute / key access, like ``pyspark.sql.Row``."""

    def __new__(cls, *values, **named):
        if named:
            row = tuple.__new__(cls, tuple(named.values()))
            row.__fields__ = list(named.keys())
        else:
            row = tuple.__new__(cls, values)
            row.__fields__ = None
        return row

    @classmethod
    def _make(cls, fields: Sequence[str], values: Sequence) -> "Row":
        row = tuple.__new__(cls, tuple(values))
        row.__fields__ = list(fields)
        return row

    def asDict(self, recursive: bool = False) -> dict:
        if not recursive:
            return dict(zip(self.__fields__ or [], self))

        def conv(v):
            if isinstance(v, Row):
                return v.asDict(True)
            if isinstance(v, list):
                return [conv(x) for x in v]
            if isinstance(v, dict):
                return {k: conv(x) for k, x in v.items()}
            return v
        return {k: conv(v) for k, v in zip(self.__fields__ or [], self)}

    def __call__(self, *values) -> "Row":
        """``Person = Row("name", "age"); Person("Alice", 11)``: a Row of names is a row factory."""
        if self.__fields__ is not None:
            raise TypeError("only a Row created from field names can be called")
        if len(values) > len(self):
            raise ValueError(f"Can not create Row with fields {tuple(self)}, expected {len(self)} values but got {values}")
        return Row._make([str(f) for f in self], values)

    def __contains__(self, item) -> bool:
        if self.__fields__ is not None:
            return item in self.__fields__
        return tuple.__contains__(self, item)

    def __reduce__(self):
        if self.__fields__ is None:
            return (Row, tuple(self))
        return (Row._make, (list(self.__fields__), tuple(self)))

    def __getattr__(self, item):
        fields = self.__dict__.get("__fields__")
        if fields and item in fields:
            return self[fields.index(item)]
        raise AttributeError(item)

    def __getitem__(self, item):
        if isinstance(item, str):
            return tuple.__getitem__(self, self.__fields__.index(item))
        return tuple.__getitem__(self, item)

    def __repr__(self):
        if self.__fields__:
            return "Row(" + ", ".join(f"{k}={v!r}" for k, v in zip(self.__fields__, self)) + ")"
        return "Row" + tuple.__repr__(self)


def _python_value(value, t: pa.DataType):
    """Arrow's ``to_pylist`` shapes -> what PySpark hands back: structs as Row, maps as dict (recursively)."""
    if value is None:
        return None
    if pa.types.is_struct(t):
        return Row._make([f.name for f in t], [_python_value(value.get(f.name), f.type) for f in t])
    if pa.types.is_map(t):
        return {_python_value(k, t.key_type): _python_value(v, t.item_type) for k, v in value}
    if pa.types.is_list(t) or pa.types.is_large_list(t) or pa.types.is_fixed_size_list(t):
        inner = t.value_type
        if _is_nested(inner):
            return [_python_value(v, inner) for v in value]
    return value


def _is_nested(t: pa.DataType) -> bool:
    return pa.types.is_struct(t) or pa.types.is_map(t) or pa.types.is_list(t) or pa.types.is_large_list(t) or \
        pa.types.is_fixed_size_list(t)


def table_to_rows(table: pa.Table) -> List["Row"]:
    names = table.column_names
    if not names:
        return []
    cols = []
    for field, col in zip(table.schema, table.columns):
        values = col.to_pylist()
        if _is_nested(field.type) and not (
                (pa.types.is_list(field.type) or pa.types.is_large_list(field.type)) and not _is_nested(field.type.value_type)):
            values = [_python_value(v, field.type) for v in values]
        cols.append(values)
    return [Row._make(names, vals) for vals in zip(*cols)]
