"""Persistence for the single-box master (sqlite3).

Parity: reference ``master/internal/db`` + ``master/static/migrations`` (Postgres via bun; 339
migrations).  The single-box master keeps the same *entities* — users/sessions, experiments
(+searcher snapshot), trials, metrics, checkpoints (v2 shape: uuid, task, allocation, resources,
metadata), allocations, task logs, templates, model registry, workspaces/projects, webhooks — in
one sqlite file with JSON columns.  Every write goes through one connection guarded by a lock;
WAL mode lets CLI reads proceed concurrently.
"""
import json
import os
import sqlite3
import threading
import time
from typing import Any, Dict, Iterable, List, Optional, Tuple

SCHEMA = """
CREATE TABLE IF NOT EXISTS meta (key TEXT PRIMARY KEY, value TEXT);
CREATE TABLE IF NOT EXISTS users (id INTEGER PRIMARY KEY AUTOINCREMENT, username TEXT UNIQUE, password_hash TEXT,
    admin INTEGER DEFAULT 0, active INTEGER DEFAULT 1, display_name TEXT, modified_at REAL);
CREATE TABLE IF NOT EXISTS sessions (token TEXT PRIMARY KEY, user_id INTEGER, expiry REAL, task_id TEXT);
CREATE TABLE IF NOT EXISTS workspaces (id INTEGER PRIMARY KEY AUTOINCREMENT, name TEXT UNIQUE, archived INTEGER DEFAULT 0, user_id INTEGER, immutable INTEGER DEFAULT 0);
CREATE TABLE IF NOT EXISTS projects (id INTEGER PRIMARY KEY AUTOINCREMENT, name TEXT, workspace_id INTEGER, description TEXT,
    archived INTEGER DEFAULT 0, user_id INTEGER, notes TEXT DEFAULT '[]', immutable INTEGER DEFAULT 0, UNIQUE(name, workspace_id));
CREATE TABLE IF NOT EXISTS experiments (id INTEGER PRIMARY KEY AUTOINCREMENT, state TEXT, config TEXT, original_config TEXT,
    model_definition BLOB, owner_id INTEGER, start_time REAL, end_time REAL, progress REAL DEFAULT 0, archived INTEGER DEFAULT 0,
    parent_id INTEGER, notes TEXT DEFAULT '', project_id INTEGER DEFAULT 1, searcher_snapshot TEXT, unmanaged INTEGER DEFAULT 0,
    job_id TEXT);
CREATE TABLE IF NOT EXISTS trials (id INTEGER PRIMARY KEY AUTOINCREMENT, experiment_id INTEGER, request_id TEXT, state TEXT,
    hparams TEXT, seed INTEGER, start_time REAL, end_time REAL, restarts INTEGER DEFAULT 0, run_id INTEGER DEFAULT 0,
    latest_checkpoint TEXT, warm_start_checkpoint TEXT, best_validation REAL, total_batches INTEGER DEFAULT 0,
    searcher_state TEXT, runner_state TEXT, summary_metrics TEXT, task_id TEXT, last_activity REAL);
CREATE INDEX IF NOT EXISTS trials_exp ON trials(experiment_id);
CREATE TABLE IF NOT EXISTS metrics (id INTEGER PRIMARY KEY AUTOINCREMENT, trial_id INTEGER, trial_run_id INTEGER, grp TEXT,
    steps_completed INTEGER, metrics TEXT, batch_metrics TEXT, report_time REAL, archived INTEGER DEFAULT 0);
CREATE INDEX IF NOT EXISTS metrics_trial ON metrics(trial_id, grp, steps_completed);
CREATE TABLE IF NOT EXISTS checkpoints (uuid TEXT PRIMARY KEY, trial_id INTEGER, task_id TEXT, allocation_id TEXT, state TEXT,
    resources TEXT, metadata TEXT, report_time REAL, steps_completed INTEGER, storage_id INTEGER, size INTEGER DEFAULT 0);
CREATE INDEX IF NOT EXISTS ckpt_trial ON checkpoints(trial_id);
CREATE TABLE IF NOT EXISTS allocations (id TEXT PRIMARY KEY, task_id TEXT, trial_id INTEGER, state TEXT, slots INTEGER,
    slot_ids TEXT, start_time REAL, end_time REAL, exit_code INTEGER, exit_reason TEXT, resource_pool TEXT);
CREATE TABLE IF NOT EXISTS tasks (id TEXT PRIMARY KEY, type TEXT, start_time REAL, end_time REAL, job_id TEXT, config TEXT, state TEXT, owner_id INTEGER);
CREATE TABLE IF NOT EXISTS task_logs (id INTEGER PRIMARY KEY AUTOINCREMENT, task_id TEXT, allocation_id TEXT, ts REAL, level TEXT,
    rank_id INTEGER, stdtype TEXT, source TEXT, log TEXT);
CREATE INDEX IF NOT EXISTS logs_task ON task_logs(task_id, id);
CREATE INDEX IF NOT EXISTS alloc_task ON allocations(task_id);
CREATE INDEX IF NOT EXISTS alloc_time ON allocations(start_time);
CREATE TABLE IF NOT EXISTS templates (name TEXT PRIMARY KEY, config TEXT, workspace_id INTEGER DEFAULT 1);
CREATE TABLE IF NOT EXISTS models (id INTEGER PRIMARY KEY AUTOINCREMENT, name TEXT UNIQUE, description TEXT, metadata TEXT DEFAULT '{}',
    labels TEXT DEFAULT '[]', user_id INTEGER, creation_time REAL, last_updated_time REAL, archived INTEGER DEFAULT 0, workspace_id INTEGER DEFAULT 1, notes TEXT DEFAULT '');
CREATE TABLE IF NOT EXISTS model_versions (id INTEGER PRIMARY KEY AUTOINCREMENT, model_id INTEGER, version INTEGER, checkpoint_uuid TEXT,
    name TEXT, comment TEXT, metadata TEXT DEFAULT '{}', labels TEXT DEFAULT '[]', user_id INTEGER, creation_time REAL, notes TEXT DEFAULT '');
CREATE TABLE IF NOT EXISTS webhooks (id INTEGER PRIMARY KEY AUTOINCREMENT, url TEXT, webhook_type TEXT, triggers TEXT);
CREATE TABLE IF NOT EXISTS trial_source_infos (id INTEGER PRIMARY KEY AUTOINCREMENT, trial_id INTEGER, checkpoint_uuid TEXT,
    type TEXT, model_id INTEGER, model_version INTEGER);
CREATE TABLE IF NOT EXISTS groups (id INTEGER PRIMARY KEY AUTOINCREMENT, name TEXT UNIQUE);
CREATE TABLE IF NOT EXISTS group_members (group_id INTEGER, user_id INTEGER, UNIQUE(group_id, user_id));
CREATE TABLE IF NOT EXISTS attrs (kind TEXT, id TEXT, key TEXT, value TEXT, PRIMARY KEY(kind, id, key));
CREATE TABLE IF NOT EXISTS rp_bindings (pool TEXT, workspace_id INTEGER, UNIQUE(pool, workspace_id));
CREATE TABLE IF NOT EXISTS oauth_clients (id TEXT PRIMARY KEY, name TEXT, domain TEXT, secret TEXT, created REAL);
CREATE TABLE IF NOT EXISTS role_assignments (id INTEGER PRIMARY KEY AUTOINCREMENT, role TEXT, user_id INTEGER, group_id INTEGER, workspace_id INTEGER);
"""


def J(x: Any) -> str:
    from determined_b200 import util

    return util.json_encode(x)


def U(s: Optional[str], default: Any = None) -> Any:
    if s is None:
        return default
    try:
        return json.loads(s)
    except (TypeError, ValueError):
        return default


class DB:
    def __init__(self, path: str) -> None:
        self.path = path
        if path != ":memory:":
            os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
        self._conn = sqlite3.connect(path, check_same_thread=False, isolation_level=None)
        self._conn.row_factory = sqlite3.Row
        self._lock = threading.RLock()
        with self._lock:
            if path != ":memory:":
                self._conn.execute("PRAGMA journal_mode=WAL")
            self._conn.execute("PRAGMA synchronous=NORMAL")
            self._conn.executescript(SCHEMA)
        self._bootstrap()

    def _bootstrap(self) -> None:
        from determined_b200.common.api import authentication as auth

        if self.one("SELECT id FROM users WHERE username='admin'") is None:
            self.exec("INSERT INTO users(username,password_hash,admin,active,modified_at) VALUES('admin',?,1,1,?)", (auth.salt_and_hash(""), time.time()))
        if self.one("SELECT id FROM users WHERE username='determined'") is None:
            self.exec("INSERT INTO users(username,password_hash,admin,active,modified_at) VALUES('determined',?,0,1,?)", (auth.salt_and_hash(""), time.time()))
        if self.one("SELECT id FROM workspaces WHERE id=1") is None:
            self.exec("INSERT INTO workspaces(id,name,immutable,user_id) VALUES(1,'Uncategorized',1,1)")
            self.exec("INSERT INTO projects(id,name,workspace_id,description,immutable,user_id) VALUES(1,'Uncategorized',1,'',1,1)")
        if self.get_meta("cluster_id") is None:
            import uuid

            self.set_meta("cluster_id", str(uuid.uuid4()))

    # -- primitives --------------------------------------------------------------------------------
    def exec(self, sql: str, args: Iterable[Any] = ()) -> int:
        with self._lock:
            cur = self._conn.execute(sql, tuple(args))
            return cur.lastrowid or 0

    def all(self, sql: str, args: Iterable[Any] = ()) -> List[sqlite3.Row]:
        with self._lock:
            return self._conn.execute(sql, tuple(args)).fetchall()

    def one(self, sql: str, args: Iterable[Any] = ()) -> Optional[sqlite3.Row]:
        with self._lock:
            return self._conn.execute(sql, tuple(args)).fetchone()

    def get_meta(self, key: str) -> Optional[str]:
        r = self.one("SELECT value FROM meta WHERE key=?", (key,))
        return None if r is None else r["value"]

    def set_meta(self, key: str, value: str) -> None:
        self.exec("INSERT INTO meta(key,value) VALUES(?,?) ON CONFLICT(key) DO UPDATE SET value=excluded.value", (key, value))

    def close(self) -> None:
        with self._lock:
            self._conn.close()

    # -- logs ----------------------------------------------------------------------------------------
    # -- task logs: here, or in ``log_backend`` (master/elastic.py) when one is configured --------------------------------
    log_backend: Any = None

    def add_logs(self, rows: List[Tuple[Any, ...]]) -> None:
        if self.log_backend is not None:
            return self.log_backend.add_logs(rows)
        with self._lock:
            self._conn.executemany(
                "INSERT INTO task_logs(task_id,allocation_id,ts,level,rank_id,stdtype,source,log) VALUES(?,?,?,?,?,?,?,?)", rows)

    def delete_logs_before(self, ts: float) -> int:
        if self.log_backend is not None:
            return self.log_backend.delete_logs_before(ts)
        with self._lock:
            cur = self._conn.execute("DELETE FROM task_logs WHERE ts < ?", (ts,))
            return int(cur.rowcount or 0)

    def delete_task_logs(self, task_id: str, before_ts: Optional[float] = None) -> int:
        if self.log_backend is not None:
            return self.log_backend.delete_task_logs(task_id, before_ts)
        with self._lock:
            if before_ts is None:
                cur = self._conn.execute("DELETE FROM task_logs WHERE task_id=?", (task_id,))
            else:
                cur = self._conn.execute("DELETE FROM task_logs WHERE task_id=? AND ts < ?", (task_id, before_ts))
            return int(cur.rowcount or 0)

    def log_field_values(self, task_id: str) -> Dict[str, List[Any]]:
        """Distinct values of the filterable columns of one task's logs."""
        if self.log_backend is not None:
            return self.log_backend.log_field_values(task_id)
        out: Dict[str, List[Any]] = {}
        for col in ("allocation_id", "rank_id", "stdtype", "source"):
            out[col] = [r[0] for r in self.all(f"SELECT DISTINCT {col} FROM task_logs WHERE task_id=? AND {col} IS NOT NULL", (task_id,))]
        return out

    def get_logs(self, task_id: str, after_id: int = 0, limit: int = 1000) -> List[Any]:
        if self.log_backend is not None:
            return self.log_backend.get_logs(task_id, after_id, limit)
        return self.all("SELECT * FROM task_logs WHERE task_id=? AND id>? ORDER BY id LIMIT ?", (task_id, after_id, limit))
