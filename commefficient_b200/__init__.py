"""commefficient_b200 -- Blackwell-native communication-efficient federated learning
(FetchSGD-style count-sketch / top-k / FedAvg) with the public surface of
pursueorigin/CommEfficient."""
__version__ = "0.1.0"
