"""Keys, defaults and environment of the ``"nebula"`` config section (reference: ``deepspeed/nebula/constants.py``).

    "nebula": {"enabled": true, "persistent_storage_path": "/foo/bar", "persistent_time_interval": 100,
               "num_of_version_in_retention": 2, "enable_nebula_load": true}

``NEBULA_<KEY>`` is the JSON key and ``NEBULA_<KEY>_DEFAULT`` its default, generated from one table."""
NEBULA = "nebula"

_FIELDS = {
    # constant stem                  json key                        default   meaning
    "ENABLED":                      ("enabled",                      False),  # use the service at all
    "LOAD_PATH":                    ("nebula_load_path",             None),   # tier-3 location preferred on load
    "ENABLE_NEBULA_LOAD":           ("enable_nebula_load",           True),   # load through the service when possible
    "PERSISTENT_STORAGE_PATH":      ("persistent_storage_path",      None),   # where versions are persisted
    "PERSISTENT_TIME_INTERVAL":     ("persistent_time_interval",     100),    # min seconds between persisted versions
    "NUM_OF_VERSION_IN_RETENTION":  ("num_of_version_in_retention",  2),      # versions kept
}
for _stem, (_key, _default) in _FIELDS.items():
    globals()[f"NEBULA_{_stem}"] = _key
    globals()[f"NEBULA_{_stem}_DEFAULT"] = _default
del _stem, _key, _default

# variables the service's runtime reads; the launcher forwards them to every node
NEBULA_EXPORT_ENVS = [
    "DLTS_JOB_ID", "DLTS_NUM_WORKER", "NEBULA_PERSISTENT_STORAGE_PATH", "NEBULA_PERSISTENT_TIME_INTERVAL", "AML_RUN_ID",
    "AZUREML_RUN_TOKEN", "AZUREML_WORKSPACE_SCOPE", "AZUREML_EXPERIMENT_SCOPE", "AZUREML_RUN_HISTORY_SERVICE_ENDPOINT",
    "AZUREML_RUN_ID", "NEBULA_MEMORY_BUFFER_SIZE", "AZUREML_PARAMETER_ITPJOB_NAME", "FC_TASKROLE_NAME", "FC_TASK_INDEX",
    "MASTER_HOST", "LOCAL_HOST", "AZUREML_BLOB_ACCOUNT_NAME", "AZUREML_BLOB_ACCOUNT_KEY",
]

# ITP pods keep their environment here
DLTS_POD_ENV_PATH = "/dlts-runtime/env/pod.env"
