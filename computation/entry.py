"""Container bootstrap (counterpart of the reference ``entry.py``): hand both callbacks to the
COINSTAC runtime."""
import coinstac

import local
import remote

coinstac.start(local.run, remote.run)
