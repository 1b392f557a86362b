// empty: stands in for the IPU SDK header of the same name (oracle/_ref build only)
