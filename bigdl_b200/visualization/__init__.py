from .summary import FileReader, FileWriter, Summary, TrainSummary, ValidationSummary, write_graph_topology

__all__ = ["FileReader", "FileWriter", "Summary", "TrainSummary", "ValidationSummary", "write_graph_topology"]
