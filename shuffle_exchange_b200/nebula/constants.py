"""Keys, defaults and environment of the ``"nebula"`` config section (reference: ``deepspeed/nebula/constants.py``).

    "nebula": {"enabled": true, "persistent_storage_path": "/foo/bar", "persistent_time_interval": 100,
               "num_of_version_in_retention": 2, "enable_nebula_load": true}
"""
NEBULA = "nebula"

# use the service at all
NEBULA_ENABLED = "enabled"
NEBULA_ENABLED_DEFAULT = False

# tier-3 location preferred on load
NEBULA_LOAD_PATH = "nebula_load_path"
NEBULA_LOAD_PATH_DEFAULT = None

# load through the service when possible
NEBULA_ENABLE_NEBULA_LOAD = "enable_nebula_load"
NEBULA_ENABLE_NEBULA_LOAD_DEFAULT = True

# where versions are persisted
NEBULA_PERSISTENT_STORAGE_PATH = "persistent_storage_path"
NEBULA_PERSISTENT_STORAGE_PATH_DEFAULT = None

# minimum seconds between persisted versions
NEBULA_PERSISTENT_TIME_INTERVAL = "persistent_time_interval"
NEBULA_PERSISTENT_TIME_INTERVAL_DEFAULT = 100

# versions kept
NEBULA_NUM_OF_VERSION_IN_RETENTION = "num_of_version_in_retention"
NEBULA_NUM_OF_VERSION_IN_RETENTION_DEFAULT = 2

# variables the service's runtime reads; the launcher forwards them to every node
NEBULA_EXPORT_ENVS = [
    "DLTS_JOB_ID", "DLTS_NUM_WORKER", "NEBULA_PERSISTENT_STORAGE_PATH", "NEBULA_PERSISTENT_TIME_INTERVAL", "AML_RUN_ID",
    "AZUREML_RUN_TOKEN", "AZUREML_WORKSPACE_SCOPE", "AZUREML_EXPERIMENT_SCOPE", "AZUREML_RUN_HISTORY_SERVICE_ENDPOINT",
    "AZUREML_RUN_ID", "NEBULA_MEMORY_BUFFER_SIZE", "AZUREML_PARAMETER_ITPJOB_NAME", "FC_TASKROLE_NAME", "FC_TASK_INDEX",
    "MASTER_HOST", "LOCAL_HOST", "AZUREML_BLOB_ACCOUNT_NAME", "AZUREML_BLOB_ACCOUNT_KEY",
]

# ITP pods keep their environment here
DLTS_POD_ENV_PATH = "/dlts-runtime/env/pod.env"
