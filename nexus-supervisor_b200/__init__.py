"""nexus-supervisor_b200 — B200-native failure classification for nexus-supervisor.

Host-side Python mirror of the one reference path this repo implements
(services/supervisor.go:137-374): event records in struct-of-arrays columns go
through libnxs_classify.so (hand-written sm_100a CUDA behind a C ABI) and come
back as packed per-record verdicts and per-job rows.  No CPU fallback exists.
"""
from . import _abi
from .classifier import (Classifier, ShardedClassifier, NxsError, JOB_ROW, FINISHED_MASK_DEFAULT, KIND, REASON, STAGE, ACTION,
                         FLAG_SUPERVISED, FLAG_FOUND, FLAG_CHECKPOINT, NO_WINNER, WORKLOAD_UNIFORM,
                         WORKLOAD_SKEWED, DEFAULT_SEED, unpack_verdict, pack_columns, encode_checkpoint_rows, shard_of, intern_kind, intern_reason, container_reason_to_event,
                         msg_template, cause_template, stage_name, action_name, reason_name)

__all__ = ["Classifier", "ShardedClassifier", "NxsError", "JOB_ROW", "FINISHED_MASK_DEFAULT", "KIND", "REASON", "STAGE", "ACTION",
           "FLAG_SUPERVISED", "FLAG_FOUND", "FLAG_CHECKPOINT", "NO_WINNER", "WORKLOAD_UNIFORM", "WORKLOAD_SKEWED",
           "DEFAULT_SEED", "unpack_verdict", "pack_columns", "encode_checkpoint_rows", "shard_of", "intern_kind", "intern_reason", "container_reason_to_event", "msg_template",
           "cause_template", "stage_name", "action_name", "reason_name", "_abi"]
