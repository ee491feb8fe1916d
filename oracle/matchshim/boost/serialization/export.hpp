// oracle/matchshim: empty stand-in for <boost/serialization/export.hpp> (test infrastructure); names come from facade.h
