"""B200-native Direct-Access (DA) ingest path of Cerebro-DS.

Drop-in for the reader interface of /root/reference/cerebro_gpdb/pg_page_reader.py and
/root/reference/cerebro_gpdb/da.py (`table_page_read`, `toast_page_read`, `input_fn`), backed by
hand-written sm_100a CUDA kernels behind a C-ABI shared library (include/cds_da.h).

Import as `cerebro_ds_kpgi_b200` (the directory name carries a hyphen; the alias module at the
repository root maps the identifier onto it).
"""
__version__ = "0.1.0"

__all__ = ["pg_page_reader", "da", "segment", "consumers", "etl", "mop", "hostnuma", "synth", "build"]
