from .federation import Federation
from .federation_client import FederationClient

__all__ = ["Federation", "FederationClient"]
