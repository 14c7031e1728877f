"""The SQL tables the path writes, byte-compatible with the reference's ORM schema.

ultrack/core/database.py:81-150 declares `nodes`, `overlaps`, `links`, `gt_nodes`,
`gt_links` through SQLAlchemy; this module issues the same DDL through the stdlib
`sqlite3` driver (SQLAlchemy is optional here) so that `Tracker.track()`, `solve()` and
the exporters of the reference open the result as their own database.  Column order,
types, primary / unique / foreign keys and the python-side defaults the ORM would have
written (`parent_id = hier_parent_id = -1`, shifts 0.0, `selected = 0`, `node_prob = -1.0`,
enum annotations 'UNKNOWN') follow `Base.metadata.create_all` on sqlite.

Only sqlite (file and shared-cache memory) is implemented; a postgresql URL raises.
"""
from __future__ import annotations

import contextlib
import itertools
import fcntl
import logging
import os
import sqlite3
from typing import Iterable, List, Optional, Sequence, Tuple

import numpy as np

LOG = logging.getLogger(__name__)

NO_PARENT = -1

_DDL = [
    """CREATE TABLE IF NOT EXISTS nodes (
	t INTEGER NOT NULL,
	id BIGINT NOT NULL,
	parent_id BIGINT,
	hier_parent_id BIGINT,
	t_node_id INTEGER,
	t_hier_id INTEGER,
	z FLOAT,
	y FLOAT,
	x FLOAT,
	z_shift FLOAT,
	y_shift FLOAT,
	x_shift FLOAT,
	area INTEGER,
	frontier FLOAT,
	height FLOAT,
	selected BOOLEAN,
	pickle BLOB,
	features BLOB,
	node_prob FLOAT,
	segm_annot VARCHAR(14),
	node_annot VARCHAR(7),
	appear_annot VARCHAR(7),
	disappear_annot VARCHAR(7),
	division_annot VARCHAR(7),
	PRIMARY KEY (t, id),
	UNIQUE (id)
)""",
    """CREATE TABLE IF NOT EXISTS overlaps (
	id INTEGER NOT NULL,
	node_id BIGINT,
	ancestor_id BIGINT,
	PRIMARY KEY (id),
	FOREIGN KEY(node_id) REFERENCES nodes (id),
	FOREIGN KEY(ancestor_id) REFERENCES nodes (id)
)""",
    """CREATE TABLE IF NOT EXISTS links (
	id INTEGER NOT NULL,
	source_id BIGINT,
	target_id BIGINT,
	weight FLOAT,
	annotation VARCHAR(7),
	PRIMARY KEY (id),
	FOREIGN KEY(source_id) REFERENCES nodes (id),
	FOREIGN KEY(target_id) REFERENCES nodes (id)
)""",
    """CREATE TABLE IF NOT EXISTS gt_nodes (
	id INTEGER NOT NULL,
	t INTEGER,
	label INTEGER,
	pickle BLOB,
	z FLOAT,
	y FLOAT,
	x FLOAT,
	PRIMARY KEY (id)
)""",
    """CREATE TABLE IF NOT EXISTS gt_links (
	id INTEGER NOT NULL,
	source_id BIGINT,
	target_id BIGINT,
	weight FLOAT,
	selected BOOLEAN,
	PRIMARY KEY (id),
	FOREIGN KEY(source_id) REFERENCES nodes (id),
	FOREIGN KEY(target_id) REFERENCES gt_nodes (id)
)""",
]

NODE_COLUMNS = ("t", "id", "parent_id", "hier_parent_id", "t_node_id", "t_hier_id", "z", "y", "x",
                "z_shift", "y_shift", "x_shift", "area", "frontier", "height", "selected", "pickle",
                "features", "node_prob", "segm_annot", "node_annot", "appear_annot", "disappear_annot",
                "division_annot")
_NODE_INSERT = f"INSERT INTO nodes ({', '.join(NODE_COLUMNS)}) VALUES ({', '.join('?' * len(NODE_COLUMNS))})"


def connect(database_path: str, timeout: float = 45.0) -> sqlite3.Connection:
    """Opens the database behind a `DataConfig.database_path` URL (dataconfig.py:87-103)."""
    if not database_path.startswith("sqlite:///"):
        raise NotImplementedError(
            f"ultrack_b200 writes sqlite databases only; got {database_path.split(':', 1)[0]}")
    target = database_path[len("sqlite:///"):]
    uri = target.startswith("file:")
    if uri:
        target = target.replace("&uri=true", "").replace("?uri=true&", "?").replace("?uri=true", "")
    return sqlite3.connect(target, timeout=timeout, uri=uri)


@contextlib.contextmanager
def write_lock(database_path: str):
    """Inter-process write lock for file-backed sqlite databases: the ranks of a multi-GPU
    run (one process per GPU) append to ONE file.  The reference serialises its workers with a
    `fasteners.InterProcessLock` file in the working directory (utils/multiprocessing.py:59-79);
    this is the same idea with `fcntl.flock` on `<database file>.lock`."""
    target = database_path[len("sqlite:///"):] if database_path.startswith("sqlite:///") else ""
    if not target or target.startswith("file:") or target == ":memory:":
        yield
        return
    fd = os.open(target + ".lock", os.O_CREAT | os.O_RDWR, 0o644)
    try:
        fcntl.flock(fd, fcntl.LOCK_EX)
        yield
    finally:
        fcntl.flock(fd, fcntl.LOCK_UN)
        os.close(fd)


# Bulk-load settings of the writer connections.  The pickle blobs are ~5 KB per node (stored
# Blosc chunks: 1 byte per bbox voxel), so a frame's insert is bound by sqlite copying blob bytes
# into overflow pages: 64 KB pages (set once, when the file is created) and no fsync per frame
# make it ~2.4x faster.  The file stays an ordinary sqlite database any reader opens.
# ULTRACK_B200_SQLITE_SAFE=1 keeps sqlite's defaults (what SQLAlchemy uses in the reference).
_FAST_SQLITE = os.environ.get("ULTRACK_B200_SQLITE_SAFE", "0") != "1"


def _bulk_pragmas(conn: sqlite3.Connection) -> None:
    if _FAST_SQLITE:
        conn.execute("PRAGMA synchronous=OFF")
        conn.execute("PRAGMA journal_mode=MEMORY")


def create_all(database_path: str) -> None:
    """Base.metadata.create_all(engine) (processing.py:459)."""
    with connect(database_path) as conn:
        if _FAST_SQLITE and conn.execute("PRAGMA page_count").fetchone()[0] <= 1:
            conn.execute("PRAGMA page_size=65536")
        for ddl in _DDL:
            conn.execute(ddl)


def clear_all_data(database_path: str) -> None:
    """ultrack/core/database.py clear_all_data: drops every table."""
    with connect(database_path) as conn:
        for table in ("gt_links", "gt_nodes", "links", "overlaps", "nodes"):
            conn.execute(f"DROP TABLE IF EXISTS {table}")


def clear_linking_data(database_path: str) -> None:
    """ultrack/core/linking/utils.py:12-21: resets the solution columns and empties `links`."""
    with connect(database_path) as conn:
        conn.execute("UPDATE nodes SET selected = 0, parent_id = ?", (NO_PARENT,))
        conn.execute("DELETE FROM links")


def maximum_time_from_database(database_path: str) -> int:
    """ultrack/core/database.py:152-162."""
    with connect(database_path) as conn:
        try:
            (max_t,) = conn.execute("SELECT max(t) FROM nodes").fetchone()
        except sqlite3.OperationalError:
            max_t = None
    LOG.info(f"Found max time = {max_t}")
    if max_t is None:
        raise ValueError(f"Dataset at {database_path} is empty.")
    return int(max_t)


def node_row(t: int, id: int, t_node_id: int, t_hier_id: int, z: int, y: int, x: int, area: int,
             frontier: float, height: float, blob: bytes, hier_parent_id: int = NO_PARENT,
             features: Optional[bytes] = None) -> Tuple:
    """One `NodeDB(...)` as _process builds it (processing.py:254-267,284) with the ORM defaults."""
    return (int(t), int(id), NO_PARENT, int(hier_parent_id), int(t_node_id), int(t_hier_id),
            float(z), float(y), float(x), 0.0, 0.0, 0.0, int(area), float(frontier), float(height), 0,
            blob, features, -1.0, "UNKNOWN", "UNKNOWN", "UNKNOWN", "UNKNOWN", "UNKNOWN")


def node_rows(t: int, ids, hier_parent_ids, t_node_ids, t_hier_ids, z, y, x, area, frontier, height, blobs,
              features) -> List[Tuple]:
    """`node_row` for a whole frame, column-wise (lists of Python scalars / buffers)."""
    n = len(ids)
    rep = itertools.repeat
    return list(zip(rep(int(t), n), ids, rep(NO_PARENT), hier_parent_ids, t_node_ids, t_hier_ids, z, y, x,
                    rep(0.0), rep(0.0), rep(0.0), area, frontier, height, rep(0), blobs, features, rep(-1.0),
                    rep("UNKNOWN"), rep("UNKNOWN"), rep("UNKNOWN"), rep("UNKNOWN"), rep("UNKNOWN")))


def insert_nodes(database_path: str, nodes: Sequence[Tuple], overlaps: Iterable[Tuple[int, int]]) -> None:
    """_insert_db (processing.py:62-87): one transaction per frame."""
    conn = connect(database_path)
    try:
        _bulk_pragmas(conn)
        with write_lock(database_path), conn:
            conn.execute("BEGIN IMMEDIATE")
            conn.executemany(_NODE_INSERT, nodes)
            conn.executemany("INSERT INTO overlaps (node_id, ancestor_id) VALUES (?, ?)", overlaps)
    finally:
        conn.close()


def insert_links(database_path: str, weight: np.ndarray, source_id: np.ndarray, target_id: np.ndarray) -> None:
    """DataFrame[weight, source_id, target_id].to_sql('links', append) (linking/processing.py:279-289).
    The reference's frame is one float64 array, so ids reach sqlite as REAL values; the BIGINT
    column affinity stores them back as integers -- the same happens here."""
    as_floats = lambda a: np.asarray(a, dtype=np.float64).tolist()      # noqa: E731  (Python floats, one C loop per column)
    rows = zip(as_floats(weight), as_floats(source_id), as_floats(target_id))
    conn = connect(database_path)
    try:
        _bulk_pragmas(conn)
        with write_lock(database_path), conn:
            conn.execute("BEGIN IMMEDIATE")
            conn.executemany("INSERT INTO links (weight, source_id, target_id) VALUES (?, ?, ?)", rows)
    finally:
        conn.close()


def insert_gt(database_path: str, gt_rows: Sequence[Tuple], weight: np.ndarray, source_id: np.ndarray,
              target_id: np.ndarray) -> None:
    """GTNodeDB rows (t, label, pickle, z, y, x) and the `gt_links` rows of one timepoint
    (core/match_gt.py:103-168; compute_spatial_neighbors appends to GTLinkDB.__tablename__)."""
    links = zip((float(w) for w in weight), (float(s) for s in source_id), (float(t) for t in target_id))
    conn = connect(database_path)
    try:
        with write_lock(database_path), conn:
            conn.execute("BEGIN IMMEDIATE")
            conn.executemany("INSERT INTO gt_nodes (t, label, pickle, z, y, x) VALUES (?, ?, ?, ?, ?, ?)", gt_rows)
            conn.executemany("INSERT INTO gt_links (weight, source_id, target_id) VALUES (?, ?, ?)", links)
    finally:
        conn.close()


def update_shifts(database_path: str, ids: np.ndarray, shifts_zyx: np.ndarray) -> None:
    """UPDATE nodes SET z_shift, y_shift, x_shift WHERE id = ... (imgproc/flow.py:740-751)."""
    rows = [(float(s[0]), float(s[1]), float(s[2]), int(i)) for i, s in zip(ids, shifts_zyx)]
    conn = connect(database_path)
    try:
        with write_lock(database_path), conn:
            conn.execute("BEGIN IMMEDIATE")
            conn.executemany("UPDATE nodes SET z_shift = ?, y_shift = ?, x_shift = ? WHERE id = ?", rows)
    finally:
        conn.close()


def load_frame_nodes(database_path: str, t: int) -> List[Tuple]:
    """SELECT id, pickle, z_shift, y_shift, x_shift FROM nodes WHERE t = ? (linking/processing.py:136-148)."""
    with connect(database_path) as conn:
        return conn.execute(
            "SELECT id, pickle, z_shift, y_shift, x_shift FROM nodes WHERE t = ? ORDER BY id", (int(t),)
        ).fetchall()
