"""Import-time check that the native async-IO engine is present and loadable."""
from ..ops.aio import AsyncIOBuilder

assert AsyncIOBuilder().is_compatible()
assert AsyncIOBuilder().load()
