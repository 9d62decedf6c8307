"""CPU restatements of the reference's Java hot loops — correctness oracles and the timed
"reference build" stand-in (the Java/Hadoop/DAAL code base cannot be built on this image)."""
