"""Command-line tools (log summaries)."""
