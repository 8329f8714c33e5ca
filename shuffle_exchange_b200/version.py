__version__ = "0.1.0"
__git_hash__ = "unknown"
__git_branch__ = "main"
