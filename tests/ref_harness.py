"""Import the reference's host code (/root/reference, or its unmodified copy under baseline/_ref) for
cross-checks, for generating golden fixtures and for the reference arm of bench.py. The GPU box has no
/root/reference: it sees the copy that `__graft_entry__.build()` / `ensure_copy()` left in baseline/_ref
(git-ignored, shipped by gpurun); without either everything here degrades to `available() == False`, the
tests that need it skip, and the committed fixtures under tests/golden/ carry what was learned.

Three packages the reference imports at package-import time are missing from the image
(SURVEY.md §0 item 11); in-memory stand-ins are installed for them:
  portkey_ai (createHeaders, PORTKEY_GATEWAY_URL), supabase (create_client, Client), and aiosqlite —
  the last as a *working* shim over stdlib sqlite3 covering the calls of src/db/local.py.
"""
from __future__ import annotations

import sqlite3
import sys
import types
from pathlib import Path

_MOUNT = Path("/root/reference")
_COPY = Path(__file__).resolve().parent.parent / "baseline" / "_ref"      # git-ignored; travels to the GPU box
REF = _MOUNT if (_MOUNT / "src" / "llm" / "base.py").exists() else _COPY


def ensure_copy() -> bool:
    """Copy the reference's Python tree to baseline/_ref (unmodified; the playground UI and VCS data are left
    out) so that the GPU box, which has no /root/reference, can run the reference's own host code in the
    reference arm of bench.py. Returns True when a usable tree exists afterwards."""
    import shutil
    if (_COPY / "src" / "llm" / "base.py").exists():
        return True
    if not (_MOUNT / "src" / "llm" / "base.py").exists():
        return False
    _COPY.parent.mkdir(parents=True, exist_ok=True)
    shutil.copytree(_MOUNT, _COPY, ignore=shutil.ignore_patterns(".git", "playground", "node_modules", "__pycache__"))
    return True


def available() -> bool:
    return (REF / "src" / "llm" / "base.py").exists()


# ---- aiosqlite shim ---------------------------------------------------------------------------
class _Cursor:
    def __init__(self, cur):
        self._cur = cur

    @property
    def rowcount(self):
        return self._cur.rowcount

    async def fetchall(self):
        return self._cur.fetchall()

    async def fetchone(self):
        return self._cur.fetchone()

    async def __aenter__(self):
        return self

    async def __aexit__(self, *a):
        return False


class _Exec:
    """Result of db.execute(...): awaitable AND usable as `async with ... as cursor`."""

    def __init__(self, conn, sql, params):
        self._conn, self._sql, self._params = conn, sql, params
        self._cursor = None

    def _run(self):
        if self._cursor is None:
            self._cursor = _Cursor(self._conn.execute(self._sql, self._params or ()))
        return self._cursor

    def __await__(self):
        async def go():
            return self._run()
        return go().__await__()

    async def __aenter__(self):
        return self._run()

    async def __aexit__(self, *a):
        return False


class _Conn:
    def __init__(self, path):
        self._path = path
        self._conn = None
        self._row_factory = None

    @property
    def row_factory(self):
        return self._row_factory

    @row_factory.setter
    def row_factory(self, f):
        self._row_factory = f
        if self._conn is not None:
            self._conn.row_factory = f

    async def __aenter__(self):
        self._conn = sqlite3.connect(self._path)
        if self._row_factory is not None:
            self._conn.row_factory = self._row_factory
        return self

    async def __aexit__(self, *a):
        self._conn.close()
        return False

    def execute(self, sql, params=None):
        return _Exec(self._conn, sql, params)

    async def commit(self):
        self._conn.commit()


def _install_stubs() -> None:
    if "aiosqlite" not in sys.modules:
        m = types.ModuleType("aiosqlite")
        m.connect = lambda path, **kw: _Conn(path)
        m.Row = sqlite3.Row
        sys.modules["aiosqlite"] = m
    if "portkey_ai" not in sys.modules:
        m = types.ModuleType("portkey_ai")
        m.PORTKEY_GATEWAY_URL = "http://portkey.invalid/v1"
        m.createHeaders = lambda **kw: {}
        sys.modules["portkey_ai"] = m
    if "supabase" not in sys.modules:
        m = types.ModuleType("supabase")

        class Client:  # noqa: D401 - placeholder type only
            pass

        def create_client(*a, **kw):
            raise RuntimeError("supabase is not available offline")
        m.Client, m.create_client = Client, create_client
        sys.modules["supabase"] = m


def load() -> bool:
    """Make `import src.llm`, `src.agents`, `src.kafka`, `src.db.local` work. False if no reference."""
    if not available():
        return False
    _install_stubs()
    if str(REF) not in sys.path:
        sys.path.append(str(REF))  # appended: this repo's own modules keep priority
    return True
