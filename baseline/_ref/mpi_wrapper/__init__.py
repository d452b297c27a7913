from .comm import Communicator
