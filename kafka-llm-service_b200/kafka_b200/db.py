"""ThreadStore — local SQLite persistence of threads and messages, plus the thread -> KV-block table.

Same file format as the reference's LocalDBClient (/root/reference/src/db/local.py): tables `threads`
and `messages` with the schema of :51-76, one JSON document per message (:156-165), so a database
written by either side (e.g. the bundled /root/reference/data/threads.db) opens with the other. The
duck-typed method set the server and the orchestrator call is mirrored (get_thread_messages :87,
add_message :145, add_messages :179, create_thread :219, thread_exists :252,
delete_thread_messages :286) — create_thread additionally accepts the `user_id` / `kafka_profile_id`
kwargs the reference's endpoint passes (/root/reference/server.py:580-584) and its own client lacks.

Two deliberate differences:
  * history order is `ORDER BY created_at ASC, rowid ASC`: the reference orders by a 1-second
    timestamp only (:101-105) and its own fixture has 22 same-second tie groups; rowid makes the order
    the insertion order by construction (bit-exact history assembly needs a total order);
  * a new table `kv_residency(thread_id, gpu, n_tokens, prefix_hash, block_ids, last_used)` records
    which thread's KV is resident where (BASELINE north_star). It is an observable cache directory,
    not a source of truth: on a miss the engine re-prefills from the message history.

stdlib `sqlite3` is used (aiosqlite is not in the image); calls are short and run inline in the
event loop exactly like the reference's effectively-synchronous Supabase client
(/root/reference/src/db/supabase.py).
"""
from __future__ import annotations

import json
import os
import sqlite3
import struct
import uuid
from datetime import datetime, timezone
from pathlib import Path
from typing import Any, Dict, Iterable, List, Optional

from .llm.types import Message

SCHEMA = """
CREATE TABLE IF NOT EXISTS threads (
    id TEXT PRIMARY KEY,
    created_at TEXT DEFAULT CURRENT_TIMESTAMP,
    metadata TEXT,
    sandbox_id TEXT
);
CREATE TABLE IF NOT EXISTS messages (
    id TEXT PRIMARY KEY,
    thread_id TEXT NOT NULL,
    message TEXT NOT NULL,
    metadata TEXT,
    created_at TEXT DEFAULT CURRENT_TIMESTAMP,
    FOREIGN KEY (thread_id) REFERENCES threads(id)
);
CREATE INDEX IF NOT EXISTS idx_messages_thread_id ON messages(thread_id);
CREATE TABLE IF NOT EXISTS kv_residency (
    thread_id TEXT PRIMARY KEY,
    gpu INTEGER NOT NULL,
    n_tokens INTEGER NOT NULL,
    prefix_hash TEXT NOT NULL,
    block_ids BLOB NOT NULL,
    last_used REAL NOT NULL
);
"""


def message_to_json(m: Message) -> str:
    """Stored document: role and content always (content may be null), other fields only when truthy
    — byte-for-byte what /root/reference/src/db/local.py:156-165 writes."""
    obj: Dict[str, Any] = {"role": m.role, "content": m.content}
    if m.name:
        obj["name"] = m.name
    if m.tool_calls:
        obj["tool_calls"] = m.tool_calls
    if m.tool_call_id:
        obj["tool_call_id"] = m.tool_call_id
    return json.dumps(obj)


def message_from_json(doc: str) -> Message:
    """Inverse, with the reference's flattening of multi-part content (local.py:124-132)."""
    d = json.loads(doc)
    content = d.get("content")
    if isinstance(content, list):
        parts = []
        for part in content:
            if isinstance(part, dict) and "text" in part:
                parts.append(part["text"])
            elif isinstance(part, str):
                parts.append(part)
        content = "\n".join(parts) if parts else None
    return Message(role=d.get("role", "user"), content=content, name=d.get("name"),
                   tool_calls=d.get("tool_calls"), tool_call_id=d.get("tool_call_id"))


class ThreadStore:
    def __init__(self, db_path: Optional[str] = None):
        self.db_path = db_path or os.environ.get("LOCAL_DB_PATH", str(Path("data") / "threads.db"))
        self._conn: Optional[sqlite3.Connection] = None

    # one connection per store (the reference opens one per call, local.py:98 — the dominant host cost
    # of its thread replay); WAL keeps readers and the writer from blocking each other
    def _db(self) -> sqlite3.Connection:
        if self._conn is None:
            if self.db_path != ":memory:":
                Path(self.db_path).parent.mkdir(parents=True, exist_ok=True)
            self._conn = sqlite3.connect(self.db_path, check_same_thread=False)
            self._conn.row_factory = sqlite3.Row
            if self.db_path != ":memory:":
                # write-ahead log + one fsync per checkpoint instead of per commit: a turn costs two commits
                # (user message, assistant message), and with the rollback journal each is two fsyncs — 0.8 ms
                # apiece, three quarters of the host time of a config-1 thread turn. Safe against a crash of
                # this process; the file stays readable by the reference's client (any SQLite >= 3.7).
                self._conn.execute("PRAGMA journal_mode=WAL")
                self._conn.execute("PRAGMA synchronous=NORMAL")
            self._conn.executescript(SCHEMA)
            self._conn.commit()
        return self._conn

    async def initialize(self) -> None:
        self._db()

    def close(self) -> None:
        if self._conn is not None:
            self._conn.close()
            self._conn = None

    # ---- messages ------------------------------------------------------------------------------
    async def get_thread_messages(self, thread_id: str, limit: Optional[int] = None,
                                  include_system: bool = True) -> List[Message]:
        q = "SELECT message FROM messages WHERE thread_id = ? ORDER BY created_at ASC, rowid ASC"
        params: List[Any] = [thread_id]
        if limit:
            q += " LIMIT ?"
            params.append(limit)
        out = []
        for row in self._db().execute(q, params):
            m = message_from_json(row["message"])
            if not include_system and m.role == "system":
                continue
            out.append(m)
        return out

    async def add_message(self, thread_id: str, message: Message,
                          metadata: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
        msg_id = str(uuid.uuid4())
        db = self._db()
        db.execute("INSERT INTO messages (id, thread_id, message, metadata) VALUES (?, ?, ?, ?)",
                   (msg_id, thread_id, message_to_json(message), json.dumps(metadata) if metadata else None))
        db.commit()
        return {"id": msg_id, "thread_id": thread_id}

    async def add_messages(self, thread_id: str, messages: List[Message]) -> List[Dict[str, Any]]:
        if not messages:
            return []
        db = self._db()
        out = []
        for m in messages:
            msg_id = str(uuid.uuid4())
            db.execute("INSERT INTO messages (id, thread_id, message) VALUES (?, ?, ?)",
                       (msg_id, thread_id, message_to_json(m)))
            out.append({"id": msg_id, "thread_id": thread_id})
        db.commit()
        return out

    async def delete_thread_messages(self, thread_id: str) -> int:
        db = self._db()
        n = db.execute("DELETE FROM messages WHERE thread_id = ?", (thread_id,)).rowcount
        db.execute("DELETE FROM kv_residency WHERE thread_id = ?", (thread_id,))
        db.commit()
        return n

    # ---- threads -------------------------------------------------------------------------------
    async def create_thread(self, thread_id: Optional[str] = None, metadata: Optional[Dict[str, Any]] = None,
                            system_message: Optional[str] = None, user_id: Optional[str] = None,
                            kafka_profile_id: Optional[str] = None, **_ignored: Any) -> Dict[str, Any]:
        thread_id = thread_id or str(uuid.uuid4())
        created_at = datetime.now(timezone.utc).replace(tzinfo=None).isoformat()
        meta = dict(metadata or {})
        if user_id:
            meta["user_id"] = user_id
        if kafka_profile_id:
            meta["kafka_profile_id"] = kafka_profile_id
        db = self._db()
        db.execute("INSERT INTO threads (id, created_at, metadata) VALUES (?, ?, ?)",
                   (thread_id, created_at, json.dumps(meta) if meta else None))
        db.commit()
        if system_message:
            await self.add_message(thread_id, Message(role="system", content=system_message))
        return {"id": thread_id, "created_at": created_at, "metadata": meta or None}

    async def thread_exists(self, thread_id: str) -> bool:
        return self._db().execute("SELECT 1 FROM threads WHERE id = ?", (thread_id,)).fetchone() is not None

    async def get_thread_metadata(self, thread_id: str) -> Optional[Dict[str, Any]]:
        row = self._db().execute("SELECT id, created_at, metadata FROM threads WHERE id = ?",
                                 (thread_id,)).fetchone()
        if row is None:
            return None
        return {"id": row["id"], "created_at": row["created_at"],
                "metadata": json.loads(row["metadata"]) if row["metadata"] else None}

    async def list_threads(self) -> List[str]:
        return [r["id"] for r in self._db().execute("SELECT id FROM threads ORDER BY id")]

    # ---- KV residency directory ----------------------------------------------------------------
    async def record_residency(self, rows: Iterable[Dict[str, Any]], gpu: int = 0) -> int:
        """Replace the directory of GPU `gpu` with a Scheduler.residency() snapshot."""
        db = self._db()
        db.execute("DELETE FROM kv_residency WHERE gpu = ?", (gpu,))
        n = 0
        for r in rows:
            if not r.get("thread_id"):
                continue  # entries without a thread hint are reachable by prefix match only
            blob = struct.pack(f"<{len(r['block_ids'])}i", *r["block_ids"])
            db.execute("INSERT OR REPLACE INTO kv_residency VALUES (?, ?, ?, ?, ?, ?)",
                       (r["thread_id"], gpu, int(r["n_tokens"]), f"{int(r['prefix_hash']):016x}", blob,
                        float(r["last_used"])))
            n += 1
        db.commit()
        return n

    async def upsert_residency(self, row: Dict[str, Any], gpu: int = 0) -> None:
        """Insert or refresh ONE thread's row (what the server does after each turn)."""
        if not row.get("thread_id"):
            return
        blob = struct.pack(f"<{len(row['block_ids'])}i", *row["block_ids"])
        db = self._db()
        db.execute("INSERT OR REPLACE INTO kv_residency VALUES (?, ?, ?, ?, ?, ?)",
                   (row["thread_id"], gpu, int(row["n_tokens"]), f"{int(row['prefix_hash']):016x}", blob,
                    float(row["last_used"])))
        db.commit()

    async def drop_residency(self, thread_id: str) -> None:
        db = self._db()
        db.execute("DELETE FROM kv_residency WHERE thread_id = ?", (thread_id,))
        db.commit()

    async def get_residency(self, thread_id: str) -> Optional[Dict[str, Any]]:
        row = self._db().execute("SELECT * FROM kv_residency WHERE thread_id = ?", (thread_id,)).fetchone()
        if row is None:
            return None
        blob = row["block_ids"]
        return {"thread_id": row["thread_id"], "gpu": row["gpu"], "n_tokens": row["n_tokens"],
                "prefix_hash": row["prefix_hash"], "block_ids": list(struct.unpack(f"<{len(blob) // 4}i", blob)),
                "last_used": row["last_used"]}
