"""Test and benchmark utilities: in-memory Kafka stub and synthetic record generators."""
